"""d3g-b200: B200-native (sm_100a) hot path of Deformable-3D-Gaussians behind the reference's Python API.

Public surface (mirrors the reference):
  rasterizer.GaussianRasterizationSettings / GaussianRasterizer   (diff_gaussian_rasterization)
  deform.DeformNetwork / DeformNetworkBaseline / DeformModel / DeformModelBaseline
  renderer.render                                                  (gaussian_renderer.render)
"""
__version__ = "0.1.0"
