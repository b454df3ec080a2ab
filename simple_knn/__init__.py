"""Import-name shim: the reference does ``from simple_knn._C import distCUDA2``
(/root/reference/scene/gaussian_model.py:20).  Putting this repository on PYTHONPATH makes that line resolve to the
B200 implementation (``langsplatv2_b200.knn``)."""
