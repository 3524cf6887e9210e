"""Role accounting of the conv-GEMM launches of one generate() (build with FTB_NVCC_EXTRA=-DFTB_GEMM_PROFILE, run with
FTB_CUDA_GRAPHS=0): CTA 0 prints, per launch, the cycles its producer / MMA / epilogue warps spent waiting.  Launch order
per generate(): token side bank+conv0, proj1+conv1, proj2+conv2, gru_ih; frame side lstm_ih, lin, postnet.bank,
postnet.proj1, postnet.proj2, post_proj."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
dev = torch.device('cuda', 0)
model = bench.build_model('fallback').to(dev)
x = bench.make_tokens(64, 200).to(dev)
for i in range(3):
    torch.cuda.synchronize()
    print(f'=== generate {i}', flush=True)
    model.generate(x)
    torch.cuda.synchronize()
