"""Drop-in for the reference `simple_knn` package (submodules/simple-knn); see `_C.distCUDA2`."""
