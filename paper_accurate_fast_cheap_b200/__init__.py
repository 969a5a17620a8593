"""paper_accurate_fast_cheap_b200 -- B200-native (sm_100a) bidirectional RWKV-6 WKV6 scan behind the
reference's operator API (wenet/rwkv_v6).  See DESIGN.md / INTEGRATION.md."""
from . import _lib
from ._lib import Wkv6Error, launch_count, load_torch_ops, set_algo  # noqa: F401
from .autograd import (BiWKV6, RUN_CUDA_RWKV6, RUN_CUDA_RWKV6_FP32, WKV6State, WKV_6, WKV_6_FP32,  # noqa: F401
                       bi_wkv6, wkv6)
from .wrappers import (RWKV_Tmix_x060c, RWKV_TmixWrapper, RWKV_TmixWrapper_bidirectional,  # noqa: F401
                       RWKV_TmixWrapper_bidirectional_direction_dropout,
                       RWKV_TmixWrapper_bidirectional_direction_dropout_both, WENET_ATTENTION_CLASSES_B200, GraphedInference, GraphedTraining,
                       install, install_ops)
from .ops import (bi_backward_raw, bi_forward_raw, combine, wkv6_backward_raw, wkv6_forward_raw)  # noqa: F401
from .convolution import ConvolutionModule, convmod_core, install_convolution  # noqa: F401

__all__ = [
    "WKV_6", "WKV_6_FP32", "RUN_CUDA_RWKV6", "RUN_CUDA_RWKV6_FP32", "WKV6State", "BiWKV6", "wkv6", "bi_wkv6",
    "wkv6_forward_raw", "wkv6_backward_raw", "bi_forward_raw", "bi_backward_raw", "combine",
    "RWKV_Tmix_x060c", "RWKV_TmixWrapper", "RWKV_TmixWrapper_bidirectional",
    "RWKV_TmixWrapper_bidirectional_direction_dropout", "RWKV_TmixWrapper_bidirectional_direction_dropout_both",
    "WENET_ATTENTION_CLASSES_B200", "install", "install_ops",
    "load_torch_ops", "launch_count", "set_algo", "Wkv6Error",
    "ConvolutionModule", "convmod_core", "install_convolution", "GraphedInference", "GraphedTraining",
]
