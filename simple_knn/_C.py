"""Drop-in for `simple_knn._C` (K/ext.cpp:15-17): distCUDA2, served by dynamic3dgaussians_b200.knn."""
from dynamic3dgaussians_b200.knn import distCUDA2, distCUDA2_dist, knn  # noqa: F401
