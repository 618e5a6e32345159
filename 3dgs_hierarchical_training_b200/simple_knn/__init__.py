"""Drop-in `simple_knn` package (reference: submodules/simple-knn); see simple_knn._C.distCUDA2."""
