"""BitNet kernel patches for transformers-style models — B200 edition.

Mirror of the reference's src/wrinklefree_inference/kernels/bitnet_patch.py (same names and call
shapes): `BitNetLinearNative` (:28-78), `pack_ternary_weights` (:81-112),
`extract_packed_weight_and_scale` (:115-162), `patch_model_with_native_kernels` (:165-210),
`benchmark_kernels` (:213-268).

Differences that are deliberate:
  * weights are packed in the kernel's BLOCKED I2_S layout (the reference file packs LSB-sequentially,
    which only its Python fallback understands) and kept resident in the b200_tile_v1 GPU layout;
  * the Sigma-a correction is applied (fused in the GEMV epilogue), so forward() returns the true
    `x @ W^T * weight_scale (+ bias)`, which is what the reference wrapper returns under its fallback;
  * no `.item()` host sync per call: the scales stay on the device.
"""
import logging
from typing import Optional, Tuple

import torch
import torch.nn as nn

logger = logging.getLogger(__name__)

try:
    from sgl_kernel.quantization import (
        bitnet_check_kernel_available,
        bitnet_gemm,
        bitnet_gemv,
        bitnet_quantize_activations,
        BITNET_BLOCK_SIZE,
    )
    NATIVE_KERNELS_AVAILABLE = bitnet_check_kernel_available()
except ImportError:
    NATIVE_KERNELS_AVAILABLE = False
    BITNET_BLOCK_SIZE = 128


class BitNetLinearNative(nn.Module):
    """Linear layer on the ternary GEMV/GEMM kernels. `packed_weight`: blocked I2_S u8 [N, K/4]."""

    def __init__(self, in_features: int, out_features: int, packed_weight: torch.Tensor, weight_scale: torch.Tensor,
                 bias: Optional[torch.Tensor] = None):
        super().__init__()
        if in_features % BITNET_BLOCK_SIZE != 0:
            raise ValueError(f"in_features ({in_features}) must be multiple of {BITNET_BLOCK_SIZE}")
        if tuple(packed_weight.shape) != (out_features, in_features // 4):
            raise ValueError("packed_weight must be [out_features, in_features/4]")
        from wf_bitnet import ops
        self.in_features = in_features
        self.out_features = out_features
        dev = packed_weight.device if packed_weight.is_cuda else torch.device("cuda")
        self.register_buffer("packed_weight", packed_weight.to(dev))
        self.register_buffer("weight_scale", weight_scale.reshape(-1)[:1].to(dev, torch.float32))
        self.register_buffer("tile_weight", ops.i2s_to_tile(self.packed_weight))
        if bias is not None:
            self.register_buffer("bias", bias.to(dev))
        else:
            self.bias = None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        from wf_bitnet import ops
        batch_shape = x.shape[:-1]
        x_flat = x.reshape(-1, self.in_features)
        if not x_flat.is_cuda:
            x_flat = x_flat.to(self.packed_weight.device)
        # per-tensor absmax like bitnet_quantize_activations(x_flat) (reference :62)
        q, act_scale, rowsum = ops.quantize_activations(x_flat, per_token=False)
        out = ops.ternary_linear(self.tile_weight, self.out_features, self.in_features, q, layout="tile", mode="linear",
                                 rowsum=rowsum, act_scale=act_scale, w_scale=self.weight_scale, out_dtype=torch.float32)
        if self.bias is not None:
            out = out + self.bias
        return out.view(*batch_shape, self.out_features)


def pack_ternary_weights(weight: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Pack ternary weights {-1,0,+1} to the kernel's blocked I2_S layout (on the GPU).
    Returns (packed uint8 [N, K/4], scale float tensor) like the reference function."""
    from wf_bitnet import ops
    out_features, in_features = weight.shape
    if in_features % 4 != 0:
        raise ValueError(f"in_features ({in_features}) must be divisible by 4")
    w = weight if weight.is_cuda else weight.to("cuda")
    packed, scale = ops.pack_ternary(w.float())
    return packed, scale.reshape(())


def extract_packed_weight_and_scale(linear: nn.Module, force_repack: bool = True):
    """Reference :115-162.  uint8 weights are taken as already packed for the kernel ([N, K/4]); HF-layout
    checkpoints ([N/4, K]) are detected by shape when the module exposes in/out_features."""
    from wf_bitnet import ops
    if not hasattr(linear, "weight"):
        return None, None, None
    weight = linear.weight.data
    if weight.dtype == torch.uint8:
        packed_weight = weight
        inf, outf = getattr(linear, "in_features", None), getattr(linear, "out_features", None)
        if inf is not None and outf is not None and tuple(weight.shape) == (outf // 4, inf) and outf // 4 != outf:
            packed_weight = ops.hf_to_i2s(weight.to("cuda"))
        if hasattr(linear, "weight_scale"):
            weight_scale = linear.weight_scale.data
        elif hasattr(linear, "scale"):
            weight_scale = linear.scale.data
        else:
            weight_scale = torch.tensor(1.0)
    elif force_repack and weight.numel() > 0:
        unique_vals = torch.unique(torch.round(weight.float()))
        is_ternary = len(unique_vals) <= 3 and unique_vals.abs().max() <= 1.0
        if not is_ternary:
            logger.warning(f"Non-ternary weights found (unique: {unique_vals.tolist()[:8]}), skipping {linear}")
            return None, None, None
        try:
            packed_weight, weight_scale = pack_ternary_weights(weight)
        except Exception as e:  # pragma: no cover
            logger.warning(f"Failed to repack weights for {linear}: {e}")
            return None, None, None
    else:
        logger.warning(f"Non-packed weights found, skipping native kernel for {linear}")
        return None, None, None
    bias = linear.bias.data if hasattr(linear, "bias") and linear.bias is not None else None
    return packed_weight, weight_scale, bias


def patch_model_with_native_kernels(model: nn.Module) -> int:
    """Replace compatible linear layers with `BitNetLinearNative`. Returns the number patched
    (0 plus a warning when the kernels are unavailable, like the reference :170-172)."""
    if not NATIVE_KERNELS_AVAILABLE:
        logger.warning("Native BitNet kernels not available, skipping patch")
        return 0
    patched = 0
    for name, module in list(model.named_modules()):
        if isinstance(module, BitNetLinearNative):
            continue
        if isinstance(module, nn.Linear) or "Linear" in type(module).__name__:
            packed_weight, weight_scale, bias = extract_packed_weight_and_scale(module)
            if packed_weight is None:
                continue
            out_features = packed_weight.shape[0]
            in_features = packed_weight.shape[1] * 4
            if in_features % BITNET_BLOCK_SIZE != 0:
                continue
            native_layer = BitNetLinearNative(in_features, out_features, packed_weight,
                                              torch.as_tensor(weight_scale, dtype=torch.float32), bias)
            parent_name = ".".join(name.split(".")[:-1])
            attr_name = name.split(".")[-1]
            parent = model.get_submodule(parent_name) if parent_name else model
            setattr(parent, attr_name, native_layer)
            patched += 1
    logger.info(f"Patched {patched} layers with native BitNet kernels")
    return patched


def benchmark_kernels(in_features: int = 2048, out_features: int = 2048, batch_size: int = 1, num_iterations: int = 100) -> dict:
    """Reference :213-268 — quant + gemv/gemm through the sgl_kernel names vs a torch matmul, timed with
    CUDA events on the current stream."""
    if not NATIVE_KERNELS_AVAILABLE:
        return {"error": "Native kernels not available"}
    dev = "cuda"
    packed_weight = torch.randint(0, 256, (out_features, in_features // 4), dtype=torch.uint8, device=dev)
    weight_scale = 0.1
    x = torch.randn(batch_size, in_features, device=dev)

    def native():
        x_int8, act_scale = bitnet_quantize_activations(x)
        if batch_size == 1:
            return bitnet_gemv(packed_weight, x_int8.squeeze(0), weight_scale * act_scale)
        return bitnet_gemm(packed_weight, x_int8, weight_scale * act_scale)

    dense = torch.randn(out_features, in_features, device=dev)

    def timed(fn):
        for _ in range(10):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(num_iterations):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / num_iterations

    native_ms = timed(native)
    torch_ms = timed(lambda: torch.matmul(x, dense.t()))
    return {"native_ms": native_ms, "torch_ms": torch_ms, "speedup": torch_ms / native_ms, "in_features": in_features,
            "out_features": out_features, "batch_size": batch_size}
