"""``simple_knn._C`` of the reference (/root/reference/submodules/simple-knn/ext.cpp:15-17 exports ``distCUDA2``)."""
from langsplatv2_b200.knn import distCUDA2  # noqa: F401
