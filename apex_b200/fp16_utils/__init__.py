"""apex_b200.fp16_utils: the names the reference exports from apex.fp16_utils
(apex/fp16_utils/__init__.py:1-19), bound to the B200 implementations."""
from .fp16util import (BN_convert_float, network_to_half, prep_param_lists, model_grads_to_master_grads,
                       master_params_to_model_params, tofp16, to_python_float, clip_grad_norm)
from .fused_weight_norm import Fused_Weight_Norm
from .fp16_optimizer import FP16_Optimizer
from .loss_scaler import LossScaler, DynamicLossScaler

__all__ = ["BN_convert_float", "network_to_half", "prep_param_lists", "model_grads_to_master_grads",
           "master_params_to_model_params", "tofp16", "to_python_float", "clip_grad_norm",
           "Fused_Weight_Norm", "FP16_Optimizer", "LossScaler", "DynamicLossScaler"]
