"""suffixarray_b200 -- B200-native truncated suffix-array build + batched substring query.

`SuffixArray` mirrors the reference's Cython class (suffix_array/suffix_array.pyx:110-267) over
ctypes; `suffix_array_cy.SuffixArray` is the same class in Cython (the reference's own host language,
`cdef extern from "sa_cuda.h"`; built by `__graft_entry__.build()`); `_lib.Index` is the thin ctypes
view of the C ABI in include/sa_cuda.h (libsa_cuda.so).
"""
from .suffix_array import SuffixArray  # noqa: F401

__all__ = ["SuffixArray"]
