"""apex_b200.fp16_utils: the names the reference exports from apex.fp16_utils
(apex/fp16_utils/__init__.py:1-19), bound to the B200 implementations."""
from . import fp16util as _util
from . import fp16_optimizer as _opt
from . import loss_scaler as _scaler
from . import fused_weight_norm as _wn

_EXPORTS = {
    _util: ("BN_convert_float", "network_to_half", "prep_param_lists", "model_grads_to_master_grads",
            "master_params_to_model_params", "tofp16", "to_python_float", "clip_grad_norm"),
    _wn: ("Fused_Weight_Norm",),
    _opt: ("FP16_Optimizer",),
    _scaler: ("LossScaler", "DynamicLossScaler"),
}
__all__ = []
for _module, _names in _EXPORTS.items():
    for _name in _names:
        globals()[_name] = getattr(_module, _name)
        __all__.append(_name)
del _module, _names, _name
