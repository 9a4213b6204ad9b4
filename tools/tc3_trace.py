#!/usr/bin/env python
"""Development probe: clock64 timeline of CTA 0 of the v3 dense TC kernel (build with EXTRA=-DNSDF_TC_TRACE).
issuer 0: (wait done, issue done) per slot-step it serves; epilogue warp 0: (wake, signalled) per slot-step."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import _capi, host  # noqa: E402

nn = host.NeuralNetwork()
assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
buf = torch.zeros(1024, dtype=torch.int64, device="cuda")
lib = _capi.lib()
lib.nsdf_debug_tc_trace.argtypes = [C.c_void_p]
lib.nsdf_debug_tc_trace.restype = None
out = torch.empty(256 ** 3, dtype=torch.float32, device="cuda")
nn.forward_lattice(256, precision=host.PREC_TC_FP16, out=out)
lib.nsdf_debug_tc_trace(buf.data_ptr())
nn.forward_lattice(256, precision=host.PREC_TC_FP16, out=out)
torch.cuda.synchronize()
lib.nsdf_debug_tc_trace(None)
t = buf.cpu().numpy()
iss, epi = t[:500], t[512:1012]
iss, epi = iss[iss > 0], epi[epi > 0]
t0 = min(iss[0], epi[0])
skip = int(sys.argv[1]) if len(sys.argv) > 1 else 100
ev = [(int(x - t0), "ISS " + ("wait_done " if i % 2 == 0 else "issued    ")) for i, x in enumerate(iss)]
ev += [(int(x - t0), "   EPI " + ("wake     " if i % 2 == 0 else "signalled")) for i, x in enumerate(epi)]
ev.sort()
prev = None
for k, (tt, name) in enumerate(ev):
    if k < skip or k > skip + 90:
        continue
    print(f"{tt:9d}  +{(tt - prev) if prev is not None else 0:5d}  {name}")
    prev = tt
d_iss = np.diff(iss)
print("issuer: mean wait->issued", d_iss[0::2][20:].mean(), " issued->next wait_done", d_iss[1::2][20:].mean())
d_epi = np.diff(epi)
print("epilogue warp 0: mean wake->signalled", d_epi[0::2][20:].mean(), " signalled->next wake", d_epi[1::2][20:].mean())
