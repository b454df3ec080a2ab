"""langsplatv2_b200 -- B200 (sm_100a) implementation of the LangSplatV2 rasterization hot path.

Public surface (mirrors the reference's operator API):
  GaussianRasterizationSettings, GaussianRasterizer   -- drop-in for diff_gaussian_rasterization
  decode_feature_map                                  -- codebook decode (+ L2 normalise)
  softmax_to_topk_soft_code, get_weights_and_indices  -- sparse-coefficient producer
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, FrameState, rasterize_forward,
                         rasterize_backward, render_decode, render_query)
from .decode import decode_feature_map
from .sparse import softmax_to_topk_soft_code, get_weights_and_indices

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "FrameState", "rasterize_forward",
           "rasterize_backward", "render_decode", "render_query", "decode_feature_map", "softmax_to_topk_soft_code", "get_weights_and_indices"]
