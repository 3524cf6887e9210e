import sys, torch
sys.path.insert(0, '.')
import bench
m = bench.build_model('fallback').to('cuda')
x = bench.make_tokens(64).to('cuda')
for _ in range(2):
    m.generate(x)
torch.cuda.synchronize()
