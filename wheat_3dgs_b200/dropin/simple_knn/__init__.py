"""Drop-in for the reference's `simple_knn` package: `from simple_knn._C import distCUDA2`
(scene/gaussian_model.py:20,148)."""
