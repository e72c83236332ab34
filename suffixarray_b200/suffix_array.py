"""``SuffixArray`` over ctypes: the host-side mirror of the reference's Cython class
(``suffix_array/suffix_array.pyx:110-267``), re-pointed at libsa_cuda.so.

The class logic lives ONCE in ``_host.SuffixArrayHost`` (shared with the Cython class
``suffix_array_cy.SuffixArray``, the reference's own host language for this layer); this module only
binds it to the ctypes ``_lib.Index``.  INTEGRATION.md shows the ``cdef extern from "sa_cuda.h"``
block a maintainer would drop into the reference's .pyx.

* ``SuffixArray(max_suffix_length=64)``                                  pyx:125
* ``SuffixArray(documents=docs, max_suffix_length=32)``                  README.md:23
* ``SuffixArray(csv_file=f, search_column=c, max_suffix_length=32)``     README.md:40-44
* ``construct_truncated_suffix_array_documents(documents)``              pyx:129-180
* ``construct_truncated_suffix_array_from_csv(filename, search_column)`` pyx:183-207
* ``query_records(substring, k=1000)``                                   pyx:209-267
* ``query_records_batch(substrings, k=1000)``  (new: the batched entry the GPU path wants)

There is no CPU fallback: without the built library or a GPU every call raises.
"""
from __future__ import annotations

from . import _lib as S
from ._host import SuffixArrayHost


class _Index(S.Index):
    """_lib.Index with the argument order _host.SuffixArrayHost uses."""

    def build_text(self, text, lowercase=False, tail=0):          # noqa: D102
        return S.Index.build_text(self, text, lowercase=lowercase, tail=tail)


class SuffixArray(SuffixArrayHost):
    _Index = _Index
    _build_csv_columns = staticmethod(S.Index.build_csv_columns)
    _build_text_multi = staticmethod(lambda indexes, text, lowercase: S.build_text_multi(indexes, text, lowercase=lowercase))

    @property
    def _idx(self):
        return self._parts[0]

    def stats(self) -> dict:
        return self._parts[0].stats()
