"""In-tree build of libnjucloud_b200.so with nvcc for sm_100a (no torch extension machinery:
the library has a plain C ABI and links only the CUDA runtime)."""
import glob
import hashlib
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIBDIR = os.path.join(_HERE, "lib")
LIBNAME = "libnjucloud_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-ffp-contract=off", "--shared",
]


def library_path():
    return os.path.join(LIBDIR, LIBNAME)


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cpp")))


def _stale(out, deps):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def _digest(deps):
    """Content hash of every source the library is built from (mtimes do not survive being copied
    to another machine; contents do)."""
    h = hashlib.sha256()
    for d in sorted(deps):
        h.update(os.path.basename(d).encode())
        with open(d, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_library(force=False, verbose=False):
    """Compile every CUDA/C++ source under csrc/ into lib/libnjucloud_b200.so (sm_100a)."""
    os.makedirs(LIBDIR, exist_ok=True)
    out = library_path()
    deps = sources() + glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(os.path.dirname(_HERE), "include", "*.h"))
    manifest = os.path.join(LIBDIR, "build_manifest.txt")
    digest = _digest(deps)
    if not force and os.path.exists(out) and os.path.exists(manifest) and open(manifest).read().strip() == digest:
        return out
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    # compile objects in parallel, then link
    objdir = os.path.join(LIBDIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for s in sources():
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if not force and not _stale(o, [s] + [d for d in deps if d.endswith((".h", ".cuh"))]):
            continue
        cmd = [nvcc] + [f for f in NVCC_FLAGS if f != "--shared"] + ["-c", s, "-o", o]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        log, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (s, log))
        if verbose and log:
            print(log)
    cmd = [nvcc, "--shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout)
    with open(manifest, "w") as f:
        f.write(digest + "\n")
    return out
