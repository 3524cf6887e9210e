import sys, torch
sys.path.insert(0, '.')
import bench
m = bench.build_model('fallback').to('cuda')
x = bench.make_tokens(64).to('cuda')
m.generate(x); torch.cuda.synchronize()
print('=== second run', flush=True)
st = m.__dict__['_accel']
for key in (('tok', 64, 200), ('frame', 64, 400)):
    prog = st['programs'][key]
    for info, step in zip(prog.info, prog.steps):
        print('---', key[0], info['label'], flush=True)
        step(); torch.cuda.synchronize()
