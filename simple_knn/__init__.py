"""Drop-in for the reference's `simple_knn` package (submodules_fsgs/simple-knn)."""
