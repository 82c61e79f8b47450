#!/usr/bin/env python
"""CIB_DEBUG_BOUNDS=1 run: trains the three model families for a few epochs with 256-byte canary bands around every engine buffer and
verifies after each fit that no kernel wrote outside its buffer; then damages one band on purpose to show that the check sees it.
(compute-sanitizer cannot be started on this GPU pool; this is the memcheck stand-in for the engine's own buffers.)"""
import ctypes as C
import json
import os
import sys

os.environ["CIB_DEBUG_BOUNDS"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from cori_intml_examples_b200.models import mnist, rpv
from cori_intml_examples_b200.ops import _lib
from cori_intml_examples_b200.utils import synthetic
from cori_intml_examples_b200.utils.np_utils import to_categorical


def canaries(be):
    bad = C.c_longlong(0)
    _lib.check(be.lib.cib_engine_check_canaries(be.engine, C.byref(bad)))
    return int(bad.value)


out = {}
x, y, _ = synthetic.rpv_like(300, 0, channels=3)
m = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=0.2, optimizer="Adam", device="cuda")
m.fit(x, y, batch_size=128, epochs=2, validation_split=0.2, verbose=0)           # partial last batch included
m.predict(x[:70], batch_size=128)
out["rpv3"] = canaries(m.backend)
x1, y1, _ = synthetic.rpv_like(130, 1, channels=1)
m1 = rpv.build_model((64, 64, 1), [8, 16, 32], [64], dropout=0.5, optimizer="Nadam", device="cuda")
m1.fit(x1, y1, batch_size=32, epochs=2, verbose=0)
out["rpv1_odd_batch"] = canaries(m1.backend)
xm, ym = synthetic.mnist_like(500, 0)
mm = mnist.build_model(32, 64, 128, 0.25, "Adadelta", device="cuda")
mm.fit(xm, to_categorical(ym, 10), batch_size=128, epochs=2, validation_split=0.17, verbose=0)
out["mnist"] = canaries(mm.backend)
xl, yl, _ = synthetic.rpv_like(64, 2, channels=1)
ml = rpv.build_legacy_model((64, 64, 1), 16, 32, 32, 32, 64, device="cuda")
ml.fit(xl, yl, batch_size=32, epochs=1, verbose=0)
out["legacy_strided"] = canaries(ml.backend)
# negative control: overwrite 4 bytes right in front of the batch buffer
p, n = C.c_void_p(), C.c_longlong(0)
_lib.check(m.backend.lib.cib_engine_buffer(m.backend.engine, 4, C.byref(p), C.byref(n)))


class _V(object):
    pass


v = _V(); v.__cuda_array_interface__ = {"shape": (1,), "typestr": "<f4", "data": (p.value - 4, False), "version": 3}
torch.as_tensor(v, device="cuda").fill_(1.0)
torch.cuda.synchronize()
out["negative_control_detected_bytes"] = canaries(m.backend)
out["negative_control_message"] = m.backend.lib.cib_last_error().decode()
out["ok"] = all(out[k] == 0 for k in ("rpv3", "rpv1_odd_batch", "mnist", "legacy_strided")) and out["negative_control_detected_bytes"] > 0
print(json.dumps(out))
sys.exit(0 if out["ok"] else 1)
