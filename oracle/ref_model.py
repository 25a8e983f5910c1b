"""TEST / BASELINE INFRASTRUCTURE ONLY (nothing in the product package imports it).

The UNMODIFIED reference `models.generator.Generator` with the out-of-scope StyleGAN2 synthesis network
replaced by a stub that returns given planes, and the decoder / beta / alpha parameters loaded from a
synthetic field (nfi_b200.synth.make_field) -- so that the reference's OWN `render` (run.py:176-350,
lifted by oracle/ref_loader.py), `sampler` closure and `TriplanarDecoder.forward` run on exactly the
synthetic workload bench.py and the parity tests use.  Used by oracle/make_golden.py-style fixture
generation, tests, and the `--impl reference` / `cpu_baseline` legs of bench.py (kind "reference")."""
import torch

import ref_loader


class PlanesStub(torch.nn.Module):
    """Stands in for models.stylegan.SynthesisNetwork (out of scope): returns the planes it holds."""

    def __init__(self, planes):
        super().__init__()
        self.planes = planes

    def forward(self, ws, **kw):
        return self.planes


def reference_generator(planes, field, scene_range, attention_values=10, use_sdf=True, trainable=False):
    """planes [B,96,Hp,Wp] (may require grad); field: dict W0,b0,W1,b1,beta,alpha (raw EqualizedLinear tensors)."""
    gen_mod = ref_loader.modules().generator
    torch.manual_seed(0)
    G = gen_mod.Generator(512, scene_range, attention_values=attention_values, use_viewdir=False,
                          use_encoder=False, disable_stylegan_noise=True, use_sdf=use_sdf).eval()
    G.synthesis_network = PlanesStub(planes)
    with torch.no_grad():
        G.decoder.net[0].weight.copy_(field["W0"]); G.decoder.net[0].bias.copy_(field["b0"])
        G.decoder.net[2].weight.copy_(field["W1"]); G.decoder.net[2].bias.copy_(field["b1"])
        if use_sdf:
            G.beta.copy_(field["beta"]); G.alpha.copy_(field["alpha"])
    return G.requires_grad_(trainable)


def reference_render_fn(script="run.py", **globals_kw):
    """The reference's `render`, bound to args / dataset_config globals (see ref_loader.lift_render)."""
    fn, _ = ref_loader.lift_render(script, **globals_kw)
    return fn
