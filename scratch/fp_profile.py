"""Three generate() calls of FastPitch at BASELINE config 3 (64 x 200, spread durations) for ncu captures
(FTB_CUDA_GRAPHS=0 so every kernel is a plain launch).  Per generate(): 20 mha_split launches (the postnet's are the last
four), 40 add_layernorm launches (the postnet's are the last eight)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
dev = torch.device('cuda', 0)
model = bench.build_model('spread', 'fast_pitch').to(dev)
x = bench.make_tokens(64, 200).to(dev)
for i in range(3):
    out = model.generate(x)
    torch.cuda.synchronize()
print('M', out['mel'].shape[2])
