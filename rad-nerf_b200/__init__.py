"""radnerf_b200 -- B200-native (sm_100a) implementation of Rad-NeRF's hot path:
occupancy-grid ray sampling -> K gated sub-NeRFs (hash-grid encode + 64-wide MLPs)
-> gated volume compositing, behind the reference's Python operator surface.

Sub-modules mirror the reference layout (thu-nics/Rad-NeRF):
  vren                      the 12 functions of models/csrc/binding.cpp:234-251
  tcnn                      Encoding / Network modules (tinycudann call sites, networks.py:42-89)
  models.custom_functions   RayAABBIntersector / RayMarcher / VolumeRenderer / TruncExp
  models.networks           NGP / MNGP / Ray_Gate
  models.rendering          render (K=1 baseline), models.ml_rendering: ml_render
  losses                    NeRFLoss
Everything computes through the C-ABI CUDA library (include/radnerf_b200.h); there is no
CPU fallback: importing an operator without the built library raises.
"""
__version__ = "0.1.0"
