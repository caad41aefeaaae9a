#!/usr/bin/env python3
"""How conflict-free the filterbank dealing of a plan came out (needs a GPU: the plan uploads its tables).
    ASSL_PLAN_DEBUG=1 python tools/plan_debug.py"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["ASSL_PLAN_DEBUG"] = "1"
from audiossl_b200 import _native  # noqa: E402

lib = _native.lib()
for n_mels, fmin, fmax in ((64, 60.0, 7800.0), (128, 0.0, 8000.0), (80, 0.0, 8000.0), (40, 20.0, 7600.0), (64, 0.0, 0.0)):
    cfg = _native.Cfg(16000, 1024, 160, n_mels, fmin, fmax)
    h = C.c_void_p()
    sys.stderr.write(f"n_mels {n_mels} fmin {fmin} fmax {fmax}: ")
    sys.stderr.flush()
    rc = lib.assl_plan_create(C.byref(h), C.byref(cfg))
    if rc:
        sys.stderr.write(f"rc {rc}\n")
    else:
        lib.assl_plan_destroy(h)
