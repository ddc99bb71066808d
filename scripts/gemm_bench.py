"""Time every GEMM shape of the MAE ViT-B step alone (CUDA events, L2 flushed between calls): TF/s and GB/s per launch.
    python scripts/gemm_bench.py [reps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4polyp_b200 import ops
from ssl4polyp_b200.ops import *

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
Me, Md = 256 * 50, 256 * 197
cases = []
groups = (("enc", Me, 768), ("dec", Md, 512))
if os.environ.get("GEMM_CASES") == "cls":          # the classifier fine-tune step: 128 images x 197 tokens, D = 768
    groups = (("enc", 128 * 197, 768),)
if os.environ.get("GEMM_CASES") == "large":        # ViT-L/16 pretraining: encoder D = 1024, decoder D = 512
    groups = (("enc", Me, 1024), ("dec", Md, 512))
for tag, M, D in groups:
    cases += [(f"{tag} qkv   NT bias", GEMM_NT, EPI_BIAS_BF16, M, 3 * D, D),
              (f"{tag} proj  NT resid", GEMM_NT, EPI_BIAS_RESIDUAL_F32, M, D, D),
              (f"{tag} fc1   NT gelu", GEMM_NT, EPI_BIAS_GELU_BF16, M, 4 * D, D),
              (f"{tag} fc2   NT resid", GEMM_NT, EPI_BIAS_RESIDUAL_F32, M, D, 4 * D),
              (f"{tag} dfc2  NN dgelu", GEMM_NN, EPI_DGELU_BF16, M, 4 * D, D),
              (f"{tag} dfc1  NN", GEMM_NN, EPI_BIAS_BF16, M, D, 4 * D),
              (f"{tag} dqkv  NN", GEMM_NN, EPI_BIAS_BF16, M, D, 3 * D),
              (f"{tag} dproj NN", GEMM_NN, EPI_BIAS_BF16, M, D, D),
              (f"{tag} wfc1  TN", GEMM_TN, EPI_PARTIAL_F32, 4 * D, D, M),
              (f"{tag} wfc2  TN", GEMM_TN, EPI_PARTIAL_F32, D, 4 * D, M),
              (f"{tag} wqkv  TN", GEMM_TN, EPI_PARTIAL_F32, 3 * D, D, M),
              (f"{tag} wproj TN", GEMM_TN, EPI_PARTIAL_F32, D, D, M)]
if os.environ.get("GEMM_CASES") != "cls":
    cases.append(("dec pred  NT f32", GEMM_NT, EPI_BIAS_F32, Md, 768, 512))
tot = 0.0
totals = {}
only = os.environ.get("GEMM_ONLY")          # e.g. GEMM_ONLY="dec fc1,dec dfc2" (ncu captures)
for name, layout, epi, M, N, K in cases:
    if only and not any(name.startswith(o) for o in only.split(",")):
        continue
    if layout == GEMM_NT:
        a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    elif layout == GEMM_NN:
        a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(K, N, device=dev).bfloat16()
    else:
        a = torch.randn(K, M, device=dev).bfloat16(); b = torch.randn(K, N, device=dev).bfloat16()
    bias = torch.randn(N, device=dev) if layout == GEMM_NT else None
    aux = None
    if epi == EPI_BIAS_RESIDUAL_F32: aux = torch.randn(M, N, device=dev)
    if epi == EPI_DGELU_BF16: aux = torch.randn(M, N, device=dev).bfloat16()
    out32 = epi in (EPI_BIAS_RESIDUAL_F32, EPI_BIAS_F32, EPI_PARTIAL_F32)
    out = torch.empty(M, N, device=dev, dtype=torch.float32 if out32 else torch.bfloat16)
    out2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if epi == EPI_BIAS_GELU_BF16 else None
    splits = ops.pick_splits(M, N, K) if layout == GEMM_TN else 1
    # GEMM_BN=0,256,192,128,64 sweeps the N tile; GEMM_CG=1,2 sweeps the CTA group (0 = the library's choice)
    for force_bn in [int(x) | (int(c) << 16) for c in os.environ.get("GEMM_CG", "0").split(",") for x in os.environ.get("GEMM_BN", "0").split(",")]:
        if (force_bn >> 16) == 2 and (force_bn & 0xffff) == 64:
            continue
        ts = []
        for _ in range(reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if layout == GEMM_TN:
                ops.gemm(a, b, layout, epi, out=out, splits=splits, force_bn=force_bn)
            else:
                ops.gemm(a, b, layout, epi, bias=bias, aux=aux, out=out, out2=out2, force_bn=force_bn)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        t = min(ts)
        byt = a.numel() * 2 + b.numel() * 2 + out.numel() * out.element_size() + (out2.numel() * 2 if out2 is not None else 0) + \
            (aux.numel() * aux.element_size() if aux is not None else 0)
        n_per_step = 12 if name.startswith("enc") else 8
        if "pred" in name: n_per_step = 1
        tot += t * n_per_step
        totals[force_bn] = totals.get(force_bn, 0.0) + t * n_per_step
        print(f"{name:18s} cg={force_bn >> 16} bn={force_bn & 0xffff:3d} M={M:6d} N={N:5d} K={K:6d} s={splits:2d}  {t:8.1f} us  {2.0 * M * N * K / t / 1e6:7.1f} TF/s  {byt / t / 1e3:6.0f} GB/s   x{n_per_step}")
from ssl4polyp_b200 import _lib as _L
print("max resident CTA pairs:", _L.load().ssl4p_gemm_max_cta_pairs(), "SMs:", torch.cuda.get_device_properties(0).multi_processor_count)
for k, v in totals.items():
    print(f"sum over the step's block GEMMs (cg={k >> 16} bn={k & 0xffff}): {v / 1e3:.2f} ms")
