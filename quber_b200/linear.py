"""nn.Linear layers of the path as tcgen05 GEMMs (quber_gemm_bias_act / quber_gemm_bias_act_x3).

The reference runs MSDeformAttn's four projections and the encoder FFN as `nn.Linear` -> cuBLAS
(OPS/modules/ms_deform_attn.py:98-124, MSMF/modeling/pixel_decoder/msdeformattn.py:116-131), in fp32
(`autocast(enabled=False)`, msdeformattn.py:314).  `PackedLinear` keeps the module's `nn.Linear` objects (so a
reference state_dict loads unchanged) and derives the GEMM operands from them per precision:

  "fp32"  fp32 accuracy on the TF32 tensor cores: split-operand 3xTF32 (the weight pre-split once into hi / lo)
  "tf32"  one TF32 product (10-bit mantissa)
  "bf16"  bf16 operands, fp32 accumulation

Several layers that read the same input (sampling_offsets + attention_weights) are concatenated into one GEMM.
The derived operands are keyed on identity + version of every source tensor, so a checkpoint load, an in-place edit or
`.to()` rebuilds them.
"""
import torch

from . import ops

PRECISIONS = ("fp32", "tf32", "bf16")


class PackedLinear:
    def __init__(self, *linears):
        self.linears = linears
        self._cache = {}

    def _sig(self):
        return tuple((p.data_ptr(), p._version) for l in self.linears for p in (l.weight, l.bias) if p is not None)

    def operands(self, precision):
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {PRECISIONS}")
        sig = self._sig()
        ent = self._cache.get(precision)
        if ent is None or ent[0] != sig:
            w = torch.cat([l.weight.detach() for l in self.linears], 0).float().contiguous()
            dev = w.device
            b = torch.cat([l.bias.detach().float() if l.bias is not None else torch.zeros(l.out_features, device=dev)
                           for l in self.linears], 0).contiguous()
            if precision == "fp32":
                hi, lo = ops.split_tf32(w)
                ent = (sig, (hi, lo, b))
            elif precision == "tf32":
                ent = (sig, (w, b))
            else:
                ent = (sig, (w.to(torch.bfloat16).contiguous(), b))
            self._cache[precision] = ent
        return ent[1]

    def __call__(self, x, precision, act="none", out_dtype=torch.float32):
        """act(x @ W.T + b): x [..., K] float32 (or bfloat16 with precision "bf16") -> [..., N] in out_dtype."""
        op = self.operands(precision)
        if precision == "fp32":
            return ops.gemm_bias_act_x3(x if x.dtype == torch.float32 else x.float(), op[0], op[1], op[2], act=act, out_dtype=out_dtype)
        if precision == "tf32":
            return ops.gemm_bias_act(x if x.dtype == torch.float32 else x.float(), op[0], op[1], act=act, out_dtype=out_dtype)
        return ops.gemm_bias_act(x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16), op[0], op[1], act=act, out_dtype=out_dtype)
