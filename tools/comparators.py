#!/usr/bin/env python
"""Kernel-to-beat comparators on the same box, same shapes, same clocks record.

The reference's own GPU kernels cannot run here (sgl_kernel / sglang are not installed and /root/reference does not
travel to the GPU box), so the stand-ins the image does carry are timed beside ours and labelled as such:
  * paged decode attention: vLLM 0.22's Triton decode kernel -- its header says it is adapted from the reference's
    python/sglang/srt/layers/attention/triton_ops/decode_attention.py -- and FlashInfer 0.6.11 BatchDecodeWithPagedKVCache
    (the reference's default backend calls FlashInfer 0.2.7, flashinfer_backend.py:524-559), page_size 1, shuffled table;
  * MLA absorbed decode: FlashInfer BatchMLAPagedAttentionWrapper (page_size 1);
  * FP8 block-scaled GEMM: vLLM cutlass_scaled_mm with 1x128 / 128x128 scales (bench.py roofline_gemm, tools/fp8_peak.py).

    FLASHINFER_WORKSPACE_BASE=/root/repo/tools/_ficache python tools/comparators.py > gpurun_out/comparators.json

FlashInfer JIT-compiles its kernels; tools/_ficache holds them prebuilt (git-ignored, travels with gpurun).
"""
import json
import os
import subprocess
import sys

os.environ.setdefault("FLASHINFER_WORKSPACE_BASE", "/root/repo/tools/_ficache")
os.environ.setdefault("FLASHINFER_CUDA_ARCH_LIST", "10.0a")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from iaas_sglang_b200 import ops  # noqa: E402

DEV = torch.device("cuda", 0)
HBM = 6551.4
try:
    HBM = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:  # noqa: BLE001
    pass


def clocks():
    try:
        return subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader",
                               "-i", "0"], capture_output=True, text=True, timeout=10).stdout.strip()
    except Exception as e:  # noqa: BLE001
        return str(e)


def time_layers(fn, n_layers, reps=6):
    """fn(l) launches on layer l's pool (pools rotate: every launch streams from HBM); ms per launch, CUDA events"""
    for l in range(n_layers):
        fn(l)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for l in range(n_layers):
            fn(l)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * n_layers)


def gqa_case(name, B, Hq, Hkv, D, ctx, shared=0, layers=3):
    g = torch.Generator(device=DEV).manual_seed(1234)
    unique = B * (ctx - shared) + shared
    perm = (torch.randperm(unique, device=DEV, generator=g) + 1).to(torch.int32)
    r2t = torch.zeros(B, ctx, dtype=torch.int32, device=DEV)
    if shared:
        r2t[:, :shared] = perm[:shared][None, :]
    r2t[:, shared:] = perm[shared:].view(B, ctx - shared)
    kb = [torch.randn(unique + 1, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
    vb = [torch.randn(unique + 1, Hkv, D, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
    q = torch.randn(B, Hq, D, device=DEV, dtype=torch.bfloat16, generator=g)
    seq = torch.full((B,), ctx, dtype=torch.int64, device=DEV)
    rpi = torch.arange(B, dtype=torch.int64, device=DEV)
    scale = D ** -0.5
    nbytes = B * ctx * Hkv * 2 * D * 2 + B * ctx * 4 + B * Hq * 2 * D * 2
    rec = {"case": name, "shape": f"B{B} Hq{Hq} Hkv{Hkv} D{D} ctx{ctx} shared{shared}", "algorithmic_bytes": nbytes}

    def add(label, ms, out=None, ref=None):
        rec[label] = {"us": round(ms * 1e3, 1), "gbs": round(nbytes / ms / 1e6, 1), "frac_of_measured_hbm": round(nbytes / ms / 1e6 / HBM, 4)}
        if out is not None and ref is not None:
            rec[label]["rel_l2_vs_ours"] = round(((out.float() - ref.float()).norm() / ref.float().norm()).item(), 6)

    # ---- ours (no KV append: the comparators do not append either) ----
    o = torch.empty(B, Hq, D, device=DEV, dtype=torch.bfloat16)
    S = ops.decode_attention_num_splits(B, Hq, Hkv, D, D, ctx, 0)
    ws = torch.empty((ops.decode_attention_workspace_bytes(B, Hq, D, S) + 3) // 4, dtype=torch.float32, device=DEV)

    def ours(l):
        ops.decode_attention(q, kb[l], vb[l], o, None, None, None, ws, r2t, rpi, seq, scale, 0.0, max_seq_len=ctx, num_splits=S)
    add("ours", time_layers(ours, layers))
    ours(0)
    ref = o.clone()

    # ---- vLLM's Triton decode kernel (adapted from the reference's triton_ops/decode_attention.py) ----
    try:
        from vllm.v1.attention.ops.triton_decode_attention import decode_attention_fwd

        o2 = torch.empty_like(o)
        lse = torch.empty(B, Hq, device=DEV, dtype=torch.float32)
        seq32 = seq.to(torch.int32)
        best = None
        for splits in (4, 8, 16):
            logits = torch.empty(B, Hq, splits, D + 1, device=DEV, dtype=torch.float32)

            def tri(l):
                decode_attention_fwd(q, kb[l].view(-1, 1, Hkv, D), vb[l].view(-1, 1, Hkv, D), o2, lse, r2t, seq32, logits, splits, scale, 1)
            try:
                ms = time_layers(tri, layers, reps=3)
            except Exception:  # noqa: BLE001
                def tri(l):  # noqa: F811
                    decode_attention_fwd(q, kb[l], vb[l], o2, lse, r2t, seq32, logits, splits, scale, 1)
                ms = time_layers(tri, layers, reps=3)
            if best is None or ms < best[0]:
                tri(0)
                best = (ms, splits, o2.clone())
        add("vllm_triton_decode (reference's Triton kernel, vLLM 0.22 copy)", best[0], best[2], ref)
        rec["vllm_triton_decode (reference's Triton kernel, vLLM 0.22 copy)"]["num_kv_splits"] = best[1]
    except Exception as e:  # noqa: BLE001
        rec["vllm_triton_decode"] = {"error": repr(e)[:300]}

    # ---- FlashInfer paged decode, page_size 1 ----
    try:
        import flashinfer

        wsb = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
        indptr = (torch.arange(B + 1, device=DEV, dtype=torch.int32) * ctx)
        last = torch.ones(B, device=DEV, dtype=torch.int32)
        for tc in (True, False):
            w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(wsb, "NHD", use_tensor_cores=tc)
            w.plan(indptr, r2t.reshape(-1).contiguous(), last, Hq, Hkv, D, 1, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)

            def fi(l):
                return w.run(q, (kb[l].view(-1, 1, Hkv, D), vb[l].view(-1, 1, Hkv, D)))
            ms = time_layers(fi, layers, reps=3)
            add(f"flashinfer_{flashinfer.__version__}_paged_decode tensor_cores={tc}", ms, fi(0), ref)
    except Exception as e:  # noqa: BLE001
        rec["flashinfer_paged_decode"] = {"error": repr(e)[:300]}
    return rec


def mla_case(name, B, H, ctx, layers=2):
    g = torch.Generator(device=DEV).manual_seed(1234)
    T = B * ctx
    r2t = (torch.randperm(T, device=DEV, generator=g) + 1).to(torch.int32).view(B, ctx)
    pool = [torch.randn(T + 1, 1, 576, device=DEV, dtype=torch.bfloat16, generator=g) for _ in range(layers)]
    q = torch.randn(B, H, 576, device=DEV, dtype=torch.bfloat16, generator=g)
    seq = torch.full((B,), ctx, dtype=torch.int64, device=DEV)
    rpi = torch.arange(B, dtype=torch.int64, device=DEV)
    scale = 192 ** -0.5
    nbytes = B * ctx * (1152 + 4) + B * H * (576 + 512) * 2
    rec = {"case": name, "shape": f"B{B} H{H} ctx{ctx} latent 576/512", "algorithmic_bytes": nbytes}
    o = torch.empty(B, H, 512, device=DEV, dtype=torch.bfloat16)
    S = ops.decode_attention_num_splits(B, H, 1, 576, 512, ctx, 0)
    ws = torch.empty((ops.decode_attention_workspace_bytes(B, H, 512, S) + 3) // 4, dtype=torch.float32, device=DEV)

    def ours(l):
        ops.decode_attention(q, pool[l], pool[l][..., :512], o, None, None, None, ws, r2t, rpi, seq, scale, 0.0, max_seq_len=ctx, num_splits=S)
    ms = time_layers(ours, layers)
    rec["ours"] = {"us": round(ms * 1e3, 1), "gbs": round(nbytes / ms / 1e6, 1), "frac_of_measured_hbm": round(nbytes / ms / 1e6 / HBM, 4)}
    ours(0)
    ref = o.clone()
    try:
        import flashinfer

        wsb = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
        w = flashinfer.mla.BatchMLAPagedAttentionWrapper(wsb, backend="fa2")
        qo_indptr = torch.arange(B + 1, device=DEV, dtype=torch.int32)
        kv_indptr = torch.arange(B + 1, device=DEV, dtype=torch.int32) * ctx
        kv_len = torch.full((B,), ctx, device=DEV, dtype=torch.int32)
        w.plan(qo_indptr, kv_indptr, r2t.reshape(-1).contiguous(), kv_len, H, 512, 64, 1, False, scale, torch.bfloat16, torch.bfloat16)
        qn, qp = q[..., :512].contiguous(), q[..., 512:].contiguous()

        def fi(l):
            return w.run(qn, qp, pool[l][:, :, :512].reshape(-1, 1, 512), pool[l][:, :, 512:].reshape(-1, 1, 64))
        # (FlashInfer wants the latent and rope parts as separate tensors: a copy of the pool, made outside the timing)
        ckv = [p_[:, :, :512].reshape(-1, 1, 512).contiguous() for p_ in pool]
        kpe = [p_[:, :, 512:].reshape(-1, 1, 64).contiguous() for p_ in pool]

        def fi2(l):
            return w.run(qn, qp, ckv[l], kpe[l])
        ms = time_layers(fi2, layers, reps=3)
        out = fi2(0)
        rec[f"flashinfer_{flashinfer.__version__}_mla fa2"] = {
            "us": round(ms * 1e3, 1), "gbs": round(nbytes / ms / 1e6, 1), "frac_of_measured_hbm": round(nbytes / ms / 1e6 / HBM, 4),
            "rel_l2_vs_ours": round(((out.float() - ref.float()).norm() / ref.float().norm()).item(), 6)}
    except Exception as e:  # noqa: BLE001
        rec["flashinfer_mla"] = {"error": repr(e)[:300]}
    return rec


def main():
    which = sys.argv[1:] or ["cfg2", "cfg3tp2", "cfg4tp8", "cfg2tp8", "cfg5tp8"]
    out = {"gpu": torch.cuda.get_device_name(0), "measured_hbm_gbs": HBM, "clocks_before": clocks(), "cases": []}
    table = {
        "cfg2": lambda: gqa_case("cfg2 (Llama-8B, 1 GPU)", 64, 32, 8, 128, 4096),
        "cfg3tp2": lambda: gqa_case("cfg3 per GPU at TP2 (Qwen-32B)", 128, 20, 4, 128, 4096),
        "cfg4tp8": lambda: gqa_case("cfg4 per GPU at TP8 (Llama-70B, shared prefix 2048)", 256, 8, 1, 128, 8192, shared=2048),
        "cfg2tp8": lambda: gqa_case("cfg2 per GPU at TP8 (batch 512)", 512, 4, 1, 128, 4096),
        "cfg5tp8": lambda: mla_case("cfg5 per GPU at TP8 (DeepSeek MLA, 16 heads)", 256, 16, 16384),
    }
    for k in which:
        try:
            rec = table[k]()
        except Exception as e:  # noqa: BLE001
            rec = {"case": k, "error": repr(e)[:400]}
        print(json.dumps(rec), file=sys.stderr, flush=True)
        out["cases"].append(rec)
        torch.cuda.empty_cache()
    out["clocks_after"] = clocks()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
