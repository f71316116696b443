"""iaas_sglang_b200 -- the B200-native (sm_100a) decode hot path behind SGLang's plugin surface.

    backend.B200AttnBackend   AttentionBackend plugin (init_forward_metadata / forward_decode / forward_extend)
    ops                       sgl_kernel-signature ops (fp8_blockwise_scaled_mm, int8_scaled_mm, awq_dequantize, ...)
    linear                    LinearMethod.apply() mirrors for FP8-block / INT8 / AWQ + row-parallel all-reduce
    tp                        one-process-per-GPU tensor parallel plumbing (torch.distributed, NCCL over NVLink)
    compat                    stand-ins for the reference structs when `sglang` is not installed

Host code is Python/PyTorch (device memory, streams, torch.distributed); every kernel is hand-written
CUDA in csrc/ behind the C ABI of include/b200_sgl.h.  No Triton, no FlashInfer, no CPU fallback.
"""
__version__ = "0.1.0"

from . import compat  # noqa: F401  (pure Python; the CUDA library is loaded lazily by ops/_lib)
