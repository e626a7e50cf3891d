#!/usr/bin/env python
"""Build the UNMODIFIED reference's own CPU (and, optionally, CUDA) code into ``oracle/_ref/``.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is imported by the product
(``detectron-pytorch_b200/``); only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may load what this script builds.

What it does (sources are read where they lie under ``/root/reference``; only *binaries* are
written, all of them below ``oracle/_ref/`` which is git-ignored but travels to the GPU box):

* the five hot-path Cython files (``libs/boxes/bbox.pyx``, ``cython_anchor.pyx``,
  ``cython_bbox_transform.pyx``, ``nms.pyx``, ``libs/nms/cpu_nms.pyx``) are cythonized at
  ``language_level=2`` after a 4-rule *alias* patch that NumPy 2 / Cython 3 force on us
  (``np.int_t -> np.int64_t``, ``dtype=np.int) -> dtype=np.int64)``, ``DTYPE = np.float ->
  np.float64``, ``np.float thresh -> double thresh``).  No arithmetic is touched.  The patched
  text only ever exists in a temp directory that is deleted after the C compiler ran.
* the pure-Python hot-path modules (``libs/boxes/anchor.py``, ``bbox_transform.py``,
  ``nms.py``, ``libs/layers/anchor.py``, ``libs/configs/config.py``, ``libs/nms/py_cpu_nms.py``)
  are cythonized *unpatched* (Cython's ``language_level=2`` understands ``xrange`` / ``print``),
  so that they exist on the GPU box as extension modules without any reference source shipping.
  ``np.float`` / ``np.int`` (removed from NumPy) are re-aliased at import time by
  ``oracle/ref_loader.py``.
* ``libs/layers/roi_align_tf/src/crop_and_resize.c`` is compiled unmodified against the
  12-line ``oracle/th_stub/TH/TH.h`` (a struct + five accessors) into ``libcrop_and_resize_ref.so``.
* with ``--cuda``: the reference's CUDA kernels (``libs/nms/nms_kernel.cu``,
  ``libs/layers/roi_align/src/cuda/roi_align_kernel.cu``,
  ``libs/layers/roi_align_tf/src/cuda/crop_and_resize_kernel.cu``,
  ``libs/layers/roi_target/src/cuda/roi_target_kernel.cu``) are compiled unmodified by
  nvcc for sm_100a into ``libref_cuda_*.so`` — a same-box GPU cross-check for the ``>``
  comparator NMS and for grid-point RoIAlign, which has no CPU code upstream.

Run:  python oracle/build_ref.py [--cuda] [--reference /root/reference]
"""
from __future__ import annotations

import argparse
import os
import re
import shutil
import subprocess
import sys
import sysconfig
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

# (reference-relative source, package below oracle/_ref, module name)
PYX = [
    ("libs/boxes/bbox.pyx", "libs/boxes", "cython_bbox"),
    ("libs/boxes/cython_anchor.pyx", "libs/boxes", "cython_anchor"),
    ("libs/boxes/cython_bbox_transform.pyx", "libs/boxes", "cython_bbox_transform"),
    ("libs/boxes/nms.pyx", "libs/boxes", "cython_nms"),
    ("libs/nms/cpu_nms.pyx", "libs/nms", "cpu_nms"),
]
PY = [
    ("libs/boxes/anchor.py", "libs/boxes", "anchor"),
    ("libs/boxes/bbox_transform.py", "libs/boxes", "bbox_transform"),
    ("libs/boxes/nms.py", "libs/boxes", "nms"),
    ("libs/layers/anchor.py", "libs/layers", "anchor"),
    ("libs/layers/box.py", "libs/layers", "box"),
    ("libs/layers/roi.py", "libs/layers", "roi"),
    ("libs/configs/config.py", "libs/configs", "config"),
    ("libs/nms/py_cpu_nms.py", "libs/nms", "py_cpu_nms"),
]

ALIAS_RULES = [
    (re.compile(r"np\.int_t"), "np.int64_t"),
    (re.compile(r"dtype=np\.int\)"), "dtype=np.int64)"),
    (re.compile(r"^DTYPE = np\.float\s*$", re.M), "DTYPE = np.float64"),
    (re.compile(r"np\.float thresh"), "double thresh"),
]


def run(cmd, **kw):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd, **kw)


def ext_suffix():
    return sysconfig.get_config_var("EXT_SUFFIX")


def compile_c(c_file, out_so, extra=()):
    import numpy as np

    inc = sysconfig.get_paths()["include"]
    run(["gcc", "-O2", "-fPIC", "-shared", "-w", "-fno-strict-aliasing",
         "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
         "-I", inc, "-I", np.get_include(), *extra, c_file, "-o", out_so])


def ensure_pkg(path):
    os.makedirs(path, exist_ok=True)
    p = OUT
    for part in os.path.relpath(path, OUT).split(os.sep):
        p = os.path.join(p, part)
        init = os.path.join(p, "__init__.py")
        if not os.path.exists(init):
            with open(init, "w") as f:
                f.write("# package marker written by oracle/build_ref.py (not reference code)\n")


def build_cython(ref):
    from Cython.Compiler.Main import compile as cy_compile, CompilationOptions

    tmp = tempfile.mkdtemp(prefix="_tmp_", dir=OUT)
    try:
        for rel, pkg, mod in PYX + PY:
            src = os.path.join(ref, rel)
            pkgdir = os.path.join(OUT, pkg)
            ensure_pkg(pkgdir)
            is_pyx = rel.endswith(".pyx")
            # Cython derives the init symbol from the file name, so stage under the module's name.
            # .pyx: alias-patched text; .py: byte-identical.  Both live only in `tmp`.
            work = os.path.join(tmp, pkg.replace("/", "_"))
            os.makedirs(work, exist_ok=True)
            staged = os.path.join(work, mod + (".pyx" if is_pyx else ".py"))
            with open(src) as f:
                text = f.read()
            if is_pyx:
                for pat, rep in ALIAS_RULES:
                    text = pat.sub(rep, text)
            with open(staged, "w") as f:
                f.write(text)
            c_file = os.path.join(work, mod + ".c")
            opts = CompilationOptions(language_level=2, output_file=c_file)
            res = cy_compile([staged], opts)
            if res.num_errors:
                raise RuntimeError("cython failed for %s" % rel)
            compile_c(c_file, os.path.join(pkgdir, mod + ext_suffix()))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    # LOG() shim (ours): the reference's libs/logs/log.py writes a log file under cfg.train_dir.
    ensure_pkg(os.path.join(OUT, "libs/logs"))
    with open(os.path.join(OUT, "libs/logs/log.py"), "w") as f:
        f.write("# shim written by oracle/build_ref.py (not reference code)\n"
                "def LOG(mssg, onscreen=True):\n    print(mssg)\n")
    # libs/layers/box.py imports its torch.utils.ffi extension at module level (`from .nms._ext import nms`);
    # that API no longer exists.  Only box.py's NumPy functions (matching_box, sample_rois, ...) are used.
    ensure_pkg(os.path.join(OUT, "libs/layers/nms/_ext"))
    with open(os.path.join(OUT, "libs/layers/nms/_ext/__init__.py"), "w") as f:
        f.write("# shim written by oracle/build_ref.py (not reference code)\nnms = None\n")


def build_crop_and_resize(ref):
    src = os.path.join(ref, "libs/layers/roi_align_tf/src/crop_and_resize.c")
    run(["gcc", "-O2", "-std=gnu99", "-fopenmp", "-fPIC", "-shared", "-w",
         "-I", os.path.join(HERE, "th_stub"), src,
         "-o", os.path.join(OUT, "libcrop_and_resize_ref.so"), "-lm"])


def build_cuda(ref):
    arch = ["-gencode", "arch=compute_100a,code=sm_100a"]
    common = ["nvcc", "-O2", "-shared", "-Xcompiler", "-fPIC", "-w", *arch]
    run(common + ["-I", os.path.join(ref, "libs/nms"),
                  os.path.join(ref, "libs/nms/nms_kernel.cu"),
                  "-o", os.path.join(OUT, "libref_cuda_nms.so")])
    d = os.path.join(ref, "libs/layers/roi_align/src/cuda")
    run(common + ["-I", d, os.path.join(d, "roi_align_kernel.cu"),
                  "-o", os.path.join(OUT, "libref_cuda_roi_align.so")])
    d = os.path.join(ref, "libs/layers/roi_align_tf/src/cuda")
    run(common + ["-I", d, os.path.join(d, "crop_and_resize_kernel.cu"),
                  "-o", os.path.join(OUT, "libref_cuda_crop_and_resize.so")])
    d = os.path.join(ref, "libs/layers/roi_target/src/cuda")
    run(common + ["-I", d, os.path.join(d, "roi_target_kernel.cu"),
                  "-o", os.path.join(OUT, "libref_cuda_roi_target.so")])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--cuda", action="store_true")
    args = ap.parse_args()
    if not os.path.isdir(args.reference):
        print("reference tree %s not present: nothing to build (prebuilt oracle/_ref is used as is)"
              % args.reference)
        return 0
    os.makedirs(OUT, exist_ok=True)
    build_cython(args.reference)
    build_crop_and_resize(args.reference)
    if args.cuda:
        build_cuda(args.reference)
    with open(os.path.join(OUT, "BUILT_FROM"), "w") as f:
        f.write("reference tree: %s\npython: %s\n" % (args.reference, sys.version.split()[0]))
    print("oracle/_ref built")
    return 0


if __name__ == "__main__":
    sys.exit(main())
