"""suffixarray_b200 -- B200-native truncated suffix-array build + batched substring query.

`SuffixArray` mirrors the reference's Cython class (suffix_array/suffix_array.pyx:110-267);
`_lib.Index` is the thin ctypes view of the C ABI in include/sa_cuda.h (libsa_cuda.so).
"""
from .suffix_array import SuffixArray  # noqa: F401

__all__ = ["SuffixArray"]
