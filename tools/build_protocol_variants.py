"""Build the two developer variants of libnfi_b200.so that tests/test_ws_protocol.py uses to reproduce the
round-1 pipeline deadlock of the field backward (see that file).  Output: nerf-from-image_b200/lib/
libnfi_b200_{newproto_delay,oldproto_delay}.so (git-ignored, travel to the GPU box with the snapshot)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "nerf-from-image_b200"))
import build as B  # noqa: E402

COMMON = ["-DNFI_WAIT_FLAG", "-DNFI_WAIT_SPINS=(1u<<18)", "-DNFI_BW_INJECT_SCATTER_DELAY=60"]
VARIANTS = {"newproto_delay": ("nfi_field_bwd.cu", COMMON),
            "oldproto_delay": ("nfi_field_bwd.cu", COMMON + ["-DNFI_BW_OLD_PROTOCOL"]),
            # forward kernel: the round-1 deferral of a tile's heads to the next ACTIVE tile (deadlocks on idle runs)
            "fwd_old_defer": ("nfi_field_fwd.cu", ["-DNFI_WAIT_FLAG", "-DNFI_WAIT_SPINS=(1u<<18)", "-DNFI_FWD_OLD_DEFER"]),
            "fwd_flag": ("nfi_field_fwd.cu", ["-DNFI_WAIT_FLAG", "-DNFI_WAIT_SPINS=(1u<<18)"])}


def build(force=False):
    B.build()
    hdrs = [os.path.join(B.CSRC, f) for f in os.listdir(B.CSRC) if f.endswith(".cuh")]
    hdrs.append(os.path.join(os.path.dirname(B.HERE), "include", "nfi_b200.h"))
    main_objs = [os.path.join(B.LIBDIR, s.replace(".cu", ".o")) for s in B.SOURCES]
    procs, relink = [], []
    for name, (SRC, flags) in VARIANTS.items():
        src = os.path.join(B.CSRC, SRC)
        lib = os.path.join(B.LIBDIR, "libnfi_b200_%s.so" % name)
        o = os.path.join(B.LIBDIR, "proto_%s.o" % name)
        if force or B._stale(o, [src, os.path.abspath(__file__)] + hdrs):
            cmd = [B._nvcc()] + B.NVCC_FLAGS + flags + ["-c", src, "-o", o]
            procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        relink.append((name, SRC, o, lib))
    for name, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode:
            raise RuntimeError("nvcc failed for %s:\n%s" % (name, out))
    for name, SRC, o, lib in relink:
        # the variant is the shipped library with ONE object replaced: relink whenever any object is newer
        objs = [m if os.path.basename(m) != SRC.replace(".cu", ".o") else o for m in main_objs]
        if force or B._stale(lib, objs):
            subprocess.check_call([B._nvcc(), "-shared", "-o", lib] + objs)
    return [os.path.join(B.LIBDIR, "libnfi_b200_%s.so" % n) for n in VARIANTS]


if __name__ == "__main__":
    print("\n".join(build(force="--force" in sys.argv)))
