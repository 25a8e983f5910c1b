"""Developer tool: build variants of libnfi_b200.so that differ in -D macros of one translation unit.

    python tools/ab_build.py nfi_field_fwd.cu g3w4:-DNFI_WS_GROUPS=3,-DNFI_WS_GROUP_WARPS=4 g1w8:...

writes nerf-from-image_b200/lib/libnfi_b200_<name>.so; select one with NFI_B200_LIB=<path>.
"""
import os
import subprocess
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "nerf-from-image_b200"))
import build as B  # noqa: E402


def main():
    B.build()
    src = sys.argv[1]
    procs = []
    for spec in sys.argv[2:]:
        name, flags = spec.split(":", 1)
        o = os.path.join(B.LIBDIR, "ab_%s_%s" % (name, src.replace(".cu", ".o")))
        cmd = [B._nvcc()] + B.NVCC_FLAGS + flags.split(",") + ["-Xptxas", "-v", "-c", os.path.join(B.CSRC, src), "-o", o]
        procs.append((name, o, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for name, o, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode:
            raise SystemExit("nvcc failed for %s:\n%s" % (name, out))
        lines = out.splitlines()
        for i, l in enumerate(lines):
            if "Used" in l and "registers" in l and i >= 2:
                print(name, lines[i - 2].split("'")[1][:60] if "'" in lines[i - 2] else "", lines[i - 1].strip(), l.split(":")[-1].strip())
        objs = [os.path.join(B.LIBDIR, s.replace(".cu", ".o")) if s != src else o for s in B.SOURCES]
        lib = os.path.join(B.LIBDIR, "libnfi_b200_%s.so" % name)
        subprocess.check_call([B._nvcc(), "-shared", "-o", lib] + objs)
        print("built", lib)


if __name__ == "__main__":
    main()
