"""Build the llb200 C-ABI shared library in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python lit-llama-ja_b200/build.py [--force]
"""
import os
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OUT = HERE / "lit_llama" / "libllb200.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def sources():
    return sorted(CSRC.glob("*.cu"))


def needs_build():
    if not OUT.exists():
        return True
    t = OUT.stat().st_mtime
    deps = list(CSRC.glob("*")) + [HERE.parent / "include" / "llb200.h"]
    return any(p.stat().st_mtime > t for p in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    objs = []
    procs = []
    for src in sources():
        obj = CSRC / (src.stem + ".o")
        cmd = [NVCC, *FLAGS, "-c", str(src), "-o", str(obj)]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for src, obj, p in procs:
        out, _ = p.communicate()
        log.append(f"== {src.name}\n{out}")
        if p.returncode != 0:
            sys.stderr.write("\n".join(log))
            raise RuntimeError(f"nvcc failed on {src.name}")
        objs.append(str(obj))
    cmd = [NVCC, "-shared", "-o", str(OUT), *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
    subprocess.run(cmd, check=True)
    (HERE / "csrc" / "build.log").write_text("\n".join(log))
    if verbose:
        print("\n".join(log))
    return OUT


def build_variant(name, defines, verbose=False):
    """Tuning build: the same sources with extra -D flags -> lit_llama/libllb200_<name>.so (select it at run time
    with LLB200_LIB=<path>; the default library is untouched)."""
    objdir = CSRC / f"_obj_{name}"
    objdir.mkdir(exist_ok=True)
    out = HERE / "lit_llama" / f"libllb200_{name}.so"
    procs = []
    for src in sources():
        obj = objdir / (src.stem + ".o")
        cmd = [NVCC, *FLAGS, *[f"-D{d}" for d in defines], "-c", str(src), "-o", str(obj)]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs, log = [], []
    for src, obj, p in procs:
        o, _ = p.communicate()
        log.append(f"== {src.name}\n{o}")
        if p.returncode != 0:
            sys.stderr.write("\n".join(log))
            raise RuntimeError(f"nvcc failed on {src.name}")
        objs.append(str(obj))
    subprocess.run([NVCC, "-shared", "-o", str(out), *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"], check=True)
    (objdir / "build.log").write_text("\n".join(log))
    if verbose:
        print("\n".join(log))
    return out


if __name__ == "__main__":
    if "--variant" in sys.argv:  # python build.py --variant xpred LLB_XPRED=1 [...]
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], [a for a in sys.argv[i + 2:] if not a.startswith("-")], verbose="-v" in sys.argv))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
