"""Per-role cycle accounting of cbhg_tail_kernel (build with FTB_NVCC_EXTRA=-DFTB_TAIL_PROFILE): config-2 shapes."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from forwardtacotron_b200 import engine as eng, runtime as rt
from forwardtacotron_b200.models import _accel
from forwardtacotron_b200.models.common_layers import CBHG
DEV = torch.device('cuda', 0)
for B, L, cin in ((64, 200, 256), (64, 400, 80)):
    m = CBHG(K=2, in_channels=cin, channels=256, proj_channels=[256, cin], num_highways=4, dropout=0.0).to(DEV).eval()
    pack = _accel.pack_cbhg(m)
    prog = eng.Program(rt.Handle.get(DEV), eng.Geometry(B, L))
    a = prog.buf('a', cin, torch.float16)
    a.normal_()
    xp = prog.buf('xp', pack.rnn.ih.n, torch.float16)
    prog.cbhg_tail(a, pack, xp)
    prog.use_graph = False
    for i in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); prog.run(); e1.record(); torch.cuda.synchronize()
        print(f'B {B} L {L} cin {cin}: {e0.elapsed_time(e1) * 1e3:.1f} us', flush=True)
