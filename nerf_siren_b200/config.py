"""Process-wide knobs of the host layer (precision of the MLP contraction, point budget per launch)."""
import os

_PRECISION = os.environ.get("NERFB_PRECISION", "auto").lower()     # "auto" | "bf16" | "fp32"
# upper bound on MLP rows per library call; bounds the activation workspace (fp32 arm: ~12 KB / row)
MAX_POINTS_FP32 = int(os.environ.get("NERFB_MAX_POINTS_FP32", 1 << 20))
MAX_POINTS_BF16 = int(os.environ.get("NERFB_MAX_POINTS_BF16", 1 << 22))


def set_precision(p: str) -> None:
    """"bf16": tcgen05 tensor-core arm (raises if the model shape is outside it); "fp32": CUDA-core arm;
    "auto": bf16 where the kernels support the model, else fp32."""
    global _PRECISION
    p = p.lower()
    if p not in ("auto", "bf16", "fp32"):
        raise ValueError(p)
    _PRECISION = p


def get_precision() -> str:
    return _PRECISION

# inference render through the single-call entry nerfb_render_fused_fwd (compositing in the MLP epilogue); 0 = the chain of
# separate library calls
RENDER_FUSED = int(os.environ.get("NERFB_RENDER_FUSED", "0")) != 0
