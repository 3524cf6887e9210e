import sys, torch
sys.path.insert(0, '.')
from forwardtacotron_b200 import engine as eng, pack as pk, runtime as rt
DEV = torch.device('cuda', 0)
h = rt.Handle.get(DEV)
def run(B, L, cin, cout, bn=None):
    gen = torch.Generator().manual_seed(1)
    prog = eng.Program(h, eng.Geometry(B, L))
    g = prog.geo
    a = prog.buf('a', cin)
    v = pk.round_tf32(torch.randn(B, L, cin, generator=gen))
    a.view(B, g.stride, cin)[:, :L] = v.to(DEV)
    w = torch.randn(cout, cin, generator=gen) / cin ** 0.5
    packed = pk.pack_linear(w.to(DEV), None, bn)
    out = prog.buf('out', cout)
    prog.gemm(a, packed, out, round_out=False)
    prog.flush()
    prog.run(); torch.cuda.synchronize()
    ref = v.double() @ pk.round_tf32(w).double().t()
    got = prog.view('out').cpu().double()
    err = (got - ref).abs().amax(dim=(0, 1))
    bad = [(i * packed.bn, float(err[i * packed.bn:(i + 1) * packed.bn].max())) for i in range((cout + packed.bn - 1) // packed.bn)]
    print(f'cin {cin} cout {cout} bn {packed.bn}: max err per n-tile', [(n0, round(e, 5)) for n0, e in bad])
run(2, 40, 256, 160, 80)
run(2, 40, 256, 400, 80)
run(2, 40, 256, 392)
run(2, 40, 392, 392)
run(2, 40, 392, 384, 96)
run(2, 40, 256, 240, 48)
run(2, 40, 256, 288, 144)
