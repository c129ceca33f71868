"""tools/wbc_probe.py -- whole-body tick (BASELINE config 4: 65,536 robots) timing for A/B builds of the library.
usage (GPU box): QRW_LIB=/path/lib.so python tools/wbc_probe.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")]
import numpy as np, torch
from qrw_b200 import _lib
_lib._SO = os.environ.get("QRW_LIB", _lib._SO)
from qrw_b200 import BatchedWbcTick, workloads as wl

B = int(os.environ.get("WBC_B", "65536"))
dev = torch.device("cuda", 0)
tick = BatchedWbcTick(B)
if os.environ.get("WBC_G"):
    _lib.check(tick._L.qrw_wbctick_set_group(tick._h, int(os.environ["WBC_G"])))
targs = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in wl.wbc_tick_inputs(B, seed=3)]
for _ in range(3):
    tick.compute(*targs)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    tick.compute(*targs)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
it = np.empty(B, np.int32)
print(f"{os.path.basename(_lib._SO)} G={os.environ.get('WBC_G', 'auto')}: {ms:.3f} ms per tick of {B} robots = {B / ms / 1e3:.1f} M ticks/s")
