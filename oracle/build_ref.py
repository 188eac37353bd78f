"""TEST INFRASTRUCTURE: builds oracle/_ref/ref_dcn*.so — the REFERENCE's own modulated deformable
convolution CUDA code, compiled UNMODIFIED for sm_100a from where it lies under /root/reference
(`detectron2/layers/csrc/deformable/deform_conv_cuda.cu`, `deform_conv_cuda_kernel.cu`, `deform_conv.h`)
together with the 2-function binding `oracle/ref_dcn_binding.cpp`.

No reference source is copied into the repository: nvcc reads the files in place and only the
binary lands in oracle/_ref/ (git-ignored, NOT gpurun-ignored, so it travels to the GPU box,
where /root/reference does not exist).  The reference's own build system (setup.py) is not run.

Only tests/, __graft_entry__.smoke() and bench.py's `reference_gpu` leg load the result — as the
checker / the "reference on the B200" timing, never as the product.

    python oracle/build_ref.py           # build if stale
    python oracle/build_ref.py --force
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("CTDET_REFERENCE_ROOT", "/root/reference")
REF_SRC = os.path.join(REF_ROOT, "detectron2", "layers", "csrc", "deformable")
OUT_DIR = os.path.join(HERE, "_ref")
MOD = "ref_dcn"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def out_path():
    return os.path.join(OUT_DIR, MOD + sysconfig.get_config_var("EXT_SUFFIX"))


def sources_present():
    return all(os.path.isfile(os.path.join(REF_SRC, f))
               for f in ("deform_conv.h", "deform_conv_cuda.cu", "deform_conv_cuda_kernel.cu"))


def build(force=False, verbose=True):
    """Returns the path of the built module, or None when the reference tree is absent (GPU box:
    the prebuilt file is used as it is)."""
    out = out_path()
    if not sources_present():
        return out if os.path.isfile(out) else None
    srcs = [os.path.join(REF_SRC, "deform_conv_cuda.cu"),
            os.path.join(REF_SRC, "deform_conv_cuda_kernel.cu"),
            os.path.join(HERE, "ref_dcn_binding.cpp")]
    deps = srcs + [os.path.join(REF_SRC, "deform_conv.h"), os.path.abspath(__file__)]
    if not force and os.path.isfile(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
        return out
    import torch
    from torch.utils import cpp_extension as C
    os.makedirs(os.path.join(OUT_DIR, "obj"), exist_ok=True)
    inc = []
    for p in C.include_paths("cuda") + [sysconfig.get_paths()["include"], REF_SRC]:
        inc += ["-I", p]
    defs = ["-DWITH_CUDA", "-DTORCH_EXTENSION_NAME=" + MOD, "-DTORCH_API_INCLUDE_EXTENSION_H",
            "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
    common = ["-O2", "-std=c++17", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
              "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
              "-D__CUDA_NO_HALF2_OPERATORS__", "-w"]
    objs, procs = [], []
    for s in srcs:
        o = os.path.join(OUT_DIR, "obj", os.path.basename(s) + ".o")
        objs.append(o)
        cmd = ["nvcc"] + ARCH + common + defs + inc + ["-x", "cu" if s.endswith(".cu") else "c++"]
        if not s.endswith(".cu"):       # host-only translation unit still goes through nvcc's host compiler
            cmd = ["nvcc"] + common + defs + inc + ["-x", "cu"] + ARCH
        cmd += ["-c", s, "-o", o]
        if verbose:
            print("[build_ref]", " ".join(cmd[:3]), "...", os.path.basename(s), flush=True)
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        log, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("compiling the reference source %s failed:\n%s" % (s, log[-4000:]))
    libs = []
    for p in C.library_paths("cuda"):
        libs += ["-L", p]
    tl = os.path.join(os.path.dirname(torch.__file__), "lib")
    link = ["nvcc"] + ARCH + ["-shared", "-o", out] + objs + libs + \
           ["-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart",
            "-Xlinker", "-rpath", "-Xlinker", tl]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("linking oracle/_ref failed:\n" + r.stdout[-2000:] + r.stderr[-2000:])
    return out


if __name__ == "__main__":
    p = build(force="--force" in sys.argv)
    print("oracle/_ref:", p if p else "reference tree absent and no prebuilt module")
