"""Placeholder kept for import compatibility.

The reference's Fused_Weight_Norm (apex/fp16_utils/fused_weight_norm.py:5-23) is
a stub whose forward and backward both raise NotImplementedError; it is not on
the gradient-step path.  The same behaviour is kept.
"""
from torch.autograd import Function


class Fused_Weight_Norm(Function):
    @staticmethod
    def forward(ctx, *args, **kwargs):
        raise NotImplementedError("Fused_Weight_Norm is a stub in the reference and here.")

    @staticmethod
    def backward(ctx, *args, **kwargs):
        raise NotImplementedError("Fused_Weight_Norm is a stub in the reference and here.")
