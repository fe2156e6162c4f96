"""Build recipe of the oracle's native pieces (test infrastructure only).

  oracle/_build/libnms_oracle.so   gcc, from oracle/nms_oracle.c (our C restatement)          -- always
  oracle/_ref/libpcdet_ref_cpu.so  g++, the reference's iou3d_cpu.cpp where it lies            -- if /root/reference exists
  oracle/_ref/libpcdet_ref_cuda.so nvcc sm_100a, the reference's iou3d_nms_kernel.cu + its host   -- if /root/reference exists
                                   side iou3d_nms.cpp (on a pointer-only at::Tensor stand-in)
  oracle/_ref/py/mmdet3d/models/…  the reference's three FCAF3D model files, byte for byte       -- if /root/reference exists
Reference sources never enter the repository: _ref/ is git-ignored build output that ships to the GPU box with the
snapshot, like the two libraries.  The staged .py files exist so that the `-m gpu` tests can run the reference's own model
code, unmodified, on the product's operator surfaces (fcaf3d_b200.compat) on a box that has no /root/reference.
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("FCAF3D_REFERENCE", "/root/reference")
REF_SRC = os.path.join(REF, "mmdet3d", "ops", "pcdet_nms", "src")
REF_PY = ("mmdet3d/models/backbones/me_resnet.py", "mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py",
          "mmdet3d/models/detectors/single_stage_sparse.py")
STAGED_PY = os.path.join(HERE, "_ref", "py")


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


def _stale(out, srcs):
    return not os.path.exists(out) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in srcs)


def build(force=False):
    built = {}
    bdir, rdir = os.path.join(HERE, "_build"), os.path.join(HERE, "_ref")
    os.makedirs(bdir, exist_ok=True)
    src = os.path.join(HERE, "nms_oracle.c")
    out = os.path.join(bdir, "libnms_oracle.so")
    if force or _stale(out, [src]):
        _run(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", out, src, "-lm"])
    built["nms_oracle"] = out
    if os.path.isdir(REF_SRC):
        os.makedirs(rdir, exist_ok=True)
        cpu_src, cuda_src = os.path.join(REF_SRC, "iou3d_cpu.cpp"), os.path.join(REF_SRC, "iou3d_nms_kernel.cu")
        out = os.path.join(rdir, "libpcdet_ref_cpu.so")
        wrap = os.path.join(HERE, "ref_wrap_cpu.cpp")
        if force or _stale(out, [wrap, cpu_src]):
            _run(["g++", "-O2", "-shared", "-fPIC", "-w", "-D__device__=", "-I", os.path.join(HERE, "ref_stubs"),
                  "-I", "/usr/local/cuda/include", '-DREF_CPU_SRC="%s"' % cpu_src, "-o", out, wrap])
        built["ref_cpu"] = out
        out = os.path.join(rdir, "libpcdet_ref_cuda.so")
        wrap = os.path.join(HERE, "ref_wrap_cuda.cu")
        host_wrap, host_src = os.path.join(HERE, "ref_wrap_nms_host.cpp"), os.path.join(REF_SRC, "iou3d_nms.cpp")
        if force or _stale(out, [wrap, cuda_src, host_wrap, host_src]):
            _run(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-w", "-shared",
                  "-Xcompiler", "-fPIC", '-DREF_CUDA_SRC="%s"' % cuda_src, '-DREF_NMS_SRC="%s"' % host_src,
                  "-I", os.path.join(HERE, "ref_stubs_host"), "-o", out, wrap, host_wrap])
        built["ref_cuda"] = out
        for rel in REF_PY:                      # unmodified copies next to the binaries (never tracked by git)
            src, dst = os.path.join(REF, rel), os.path.join(STAGED_PY, rel)
            if force or _stale(dst, [src]):
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                shutil.copyfile(src, dst)
        built["ref_py"] = STAGED_PY
    else:
        for name in ("libpcdet_ref_cpu.so", "libpcdet_ref_cuda.so"):
            p = os.path.join(rdir, name)
            if os.path.exists(p):
                built["ref_cpu" if "cpu" in name else "ref_cuda"] = p
        if all(os.path.exists(os.path.join(STAGED_PY, rel)) for rel in REF_PY):
            built["ref_py"] = STAGED_PY
    return built


if __name__ == "__main__":
    print(build(force=True))
