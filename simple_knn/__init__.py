"""Import-name shim for the reference's optional third-party package `simple_knn` (gaussian_model/gaussian_model.py:7-11):
`from simple_knn._C import distCUDA2` resolves to the hash-grid k-NN kernel of libnlosb200.so."""
