"""Import-name shim: the reference's extension is installed as `nlos_gaussian_renderer` with a compiled submodule `_C`
(submodules/cuda_renderer/setup.py:9-12; imported at gaussian_model/cuda_autograd.py:11, rendering_cuda.py:11,
test_cuda_autograd.py:20).  Here `_C` is the ctypes-backed mirror over libnlosb200.so; importing it fails with
ImportError when the library cannot be loaded, which is the condition the reference's callers test for."""
