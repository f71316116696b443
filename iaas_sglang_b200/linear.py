"""LinearMethod.apply() mirrors for the quantised linears of the decode path.

Reference call sites:
  Fp8LinearMethod.apply (block_quant branch)  python/sglang/srt/layers/quantization/fp8.py:418-455
     -> cutlass_w8a8_block_fp8_linear_with_fallback  fp8_utils.py:180-210
  W8A8Int8LinearMethod.apply                   python/sglang/srt/layers/quantization/w8a8_int8.py:125-146
  AWQLinearMethod.apply                        python/sglang/srt/layers/quantization/awq.py:186-204
  RowParallelLinear.forward (+ all-reduce)     python/sglang/srt/layers/linear.py:1360-1382
`layer` is any object carrying the parameter names the reference methods create.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops, tp


class Fp8BlockLinearMethod:
    """layer.weight [N,K] float8_e4m3fn, layer.weight_scale_inv [N/128, K/128] float32 (fp8.py:236-300)."""

    block_size = (128, 128)

    def apply(self, layer, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        weight, weight_scale = layer.weight, layer.weight_scale_inv
        x2d = x.view(-1, x.shape[-1])
        out_shape = [*x.shape[:-1], weight.shape[0]]
        q, xs = ops.per_token_group_quant_fp8(x2d, 128, column_major_scales=True)      # fp8_utils.py:202-204
        out = ops.fp8_blockwise_scaled_mm(q, weight.t(), xs, weight_scale.t(), x.dtype)  # fp8_utils.py:205-207
        if bias is not None:
            out = out + bias
        return out.to(dtype=x.dtype).view(*out_shape)

    def apply_quantized(self, layer, x_q: torch.Tensor, x_s: torch.Tensor, out_dtype: torch.dtype,
                        bias: Optional[torch.Tensor] = None, rope: Optional[tuple] = None) -> torch.Tensor:
        """The same linear when the producer of x already emitted its group-128 quantisation
        (ops.rmsnorm_quant_fp8 / ops.silu_and_mul_quant_fp8): fp8_utils.py:205-207 without :202-204.
        rope = (positions, rope_cols, head_dim, theta, cos_sin) for a qkv projection: the rotary embedding that follows it
        (models/llama.py:186-188) runs in the GEMM's epilogue where it can (ops.fp8_blockwise_scaled_mm)."""
        if bias is not None and rope is not None:
            raise NotImplementedError("rope fusion with a bias (the bias would have to be added before the rotation)")
        out = ops.fp8_blockwise_scaled_mm(x_q, layer.weight.t(), x_s, layer.weight_scale_inv.t(), out_dtype, rope=rope)
        return out if bias is None else out + bias


class W8A8Int8LinearMethod:
    """layer.weight stored as W^T view ([K,N] column-major int8), layer.weight_scale [N,1] f32 (w8a8_int8.py:88-123)."""

    def apply(self, layer, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        x2d = x.view(-1, x.shape[-1])
        x_q, x_scale = ops.per_token_quant_int8(x2d)                                     # w8a8_int8.py:141
        out = ops.int8_scaled_mm(x_q, layer.weight, x_scale, layer.weight_scale, out_dtype=x.dtype, bias=bias)
        return out.view(*x.shape[:-1], out.shape[-1])


class AWQLinearMethod:
    """layer.qweight [K, N/8] int32, layer.qzeros [K/g, N/8] int32, layer.scales [K/g, N] half (awq.py:104-184)."""

    def apply(self, layer, x: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        qweight, scales, qzeros = layer.qweight, layer.scales, layer.qzeros
        out_shape = x.shape[:-1] + (qweight.shape[-1] * 8,)
        out = ops.awq_w4a16_gemm(x.reshape(-1, x.shape[-1]), qweight, scales, qzeros, bias)
        return out.reshape(out_shape)


def row_parallel_forward(method, layer, x, bias: Optional[torch.Tensor] = None,
                         reduce_results: bool = True, out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """RowParallelLinear.forward (linear.py:1360-1382): bias only on rank 0, all-reduce when tp_size > 1.
    x may be an (x_q, x_s) pair from a fused quantising producer (FP8 block method only)."""
    b = bias if (bias is not None and tp.get_tensor_model_parallel_rank() == 0) else None
    if isinstance(x, tuple):
        out = method.apply_quantized(layer, x[0], x[1], out_dtype, b)
    else:
        out = method.apply(layer, x, b)
    if reduce_results and tp.get_tensor_model_parallel_world_size() > 1:
        out = tp.tensor_model_parallel_all_reduce(out)
    return out
