"""Python face of the sm_100a kernels. Import is cheap; the shared library loads on first use."""
from . import native
from .attention import flash_attn
from .elementwise import (binary, cfg_step, finalize_u8, lincomb, matting_composite, philox_normal, philox_normal_cpu,
                          philox_u32, resize_bilinear, timestep_embedding, upsample_nearest2x)
from .gemm import (DirectConvWeight, auto_splitk, batch_scale, get_batch_scale, conv3x3, conv3x3_strided, conv3x3_strided_gather, conv_direct, conv_temporal, gemm,
                   pack_conv_weight, pack_geglu_weight, pack_temporal_conv_weight)
from .temporal import temporal_attn
from .norm import groupnorm, layernorm
