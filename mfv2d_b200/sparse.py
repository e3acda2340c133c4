"""``SparseVector`` and ``MatrixCRS`` of ``mfv2d._mfv2d`` as host containers.

They are the construction path of the constraint blocks of a ``LinearSystem``
(``MatrixCRS(rows, cols).build_row(i, SparseVector.from_entries(n, idx, val))``,
python/mfv2d/solving.py:62-90) and never touch the device: what the library receives are their flat
``(row, column, value)`` triples.  Storage is three numpy arrays; method names, argument meaning, returned
dtypes and the error classes follow src/algebra/svector.c and src/algebra/crs_matrix.c (cited per method).
Checked against the compiled reference in tests/test_sparse.py.
"""
from __future__ import annotations

from typing import Iterable

import numpy as np

__all__ = ["SparseVector", "MatrixCRS"]


class SparseVector:
    """Vector which stores only the non-zero components (svector.h:13-33): sorted ``uint64`` indices and
    ``float64`` values, dimension ``n``."""

    __slots__ = ("_n", "_idx", "_val")

    def __init__(self, *args, **kwargs):
        # the reference type has no public constructor: instances come from the class methods
        if args or kwargs:
            raise TypeError("cannot create 'mfv2d._mfv2d.SparseVector' instances")
        self._n, self._idx, self._val = 0, np.zeros(0, np.uint64), np.zeros(0, np.float64)

    @classmethod
    def _make(cls, n: int, idx: np.ndarray, val: np.ndarray) -> "SparseVector":
        self = object.__new__(cls)
        self._n = int(n)
        self._idx = np.ascontiguousarray(idx, np.uint64)
        self._val = np.ascontiguousarray(val, np.float64)
        return self

    # ---- constructors ----
    @classmethod
    def from_entries(cls, n, indices, values, *extra, **kwargs) -> "SparseVector":
        """svec_from_entries, svector.c:291-433: both arrays 1-D of equal length, indices strictly
        increasing and below ``n``."""
        if kwargs:
            raise TypeError("SparseVector.from_entries() takes no keyword arguments.")
        if extra:
            raise TypeError(f"SparseVector.from_entries() takes exactly 3 arguments, got {3 + len(extra)}.")
        n = int(n)
        if n <= 0:
            raise ValueError("The dimension of the vector must be strictly positive.")
        idx = np.asarray(indices)
        if idx.size == 0:
            idx = np.zeros(idx.shape, np.uint64)
        elif not np.issubdtype(idx.dtype, np.integer):
            raise TypeError("indices can not be safely converted to unsigned integers")
        elif np.any(idx < 0):
            raise OverflowError("indices can not be negative")
        idx = idx.astype(np.uint64)
        val = np.asarray(values, dtype=np.float64)
        if idx.ndim != 1 or val.ndim != 1:
            raise ValueError("indices and values must be one dimensional arrays")
        if idx.size != val.size:
            raise ValueError(
                "Indices and values must be arrays of equal length. Instead indices had "
                f"{idx.size} elements and values had {val.size} elements."
            )
        bad = np.flatnonzero(idx[:-1] >= idx[1:])
        if bad.size:
            raise ValueError(f"Entry indices at {bad[0] + 1} and at {bad[0] + 2} are not sorted.")
        out = np.flatnonzero(idx >= np.uint64(n))
        if out.size:
            raise ValueError(f"Entry index at {out[0]} with value {idx[out[0]]} is outside the allowed range [0, {n}).")
        return cls._make(n, idx.copy(), val.copy())

    @classmethod
    def from_pairs(cls, *args, **kwargs) -> "SparseVector":
        """svec_from_pairs, svector.c:674-755: ``(n, (index, value), ...)``; pairs are sorted by index,
        duplicates are an error."""
        if kwargs:
            raise TypeError("SparseVector.from_pairs() takes no keyword arguments.")
        if len(args) < 2:
            raise TypeError("At least one argument must be provided." if not args else "At least one pair must be provided.")
        n = int(args[0])
        if n <= 0:
            raise ValueError("Vector dimension count must be positive.")
        idx = np.empty(len(args) - 1, np.uint64)
        val = np.empty(len(args) - 1, np.float64)
        for k, pair in enumerate(args[1:]):
            if not isinstance(pair, tuple) or len(pair) != 2:
                raise TypeError(f"Each pair must be a tuple (int, float), got {pair!r}")
            i, v = int(pair[0]), float(pair[1])
            if i < 0:
                raise ValueError("DoF index must be non-negative.")
            if not np.isfinite(v):
                raise ValueError("Coefficient must be finite.")
            if i >= n:
                raise ValueError(f"DoF index {i} is outside the allowed range [0, {n}).")
            idx[k], val[k] = i, v
        order = np.argsort(idx, kind="stable")
        idx, val = idx[order], val[order]
        if np.any(idx[:-1] == idx[1:]):
            raise ValueError("Duplicate DoF indices in the pairs.")
        return cls._make(n, idx, val)

    @classmethod
    def concatenate(cls, *vectors: "SparseVector", **kwargs) -> "SparseVector":
        """svec_concatenate, svector.c:623-672: dimensions add up, indices are shifted."""
        if kwargs:
            raise TypeError("SparseVector.concatenate() takes no keyword arguments.")
        shift, idx, val = 0, [], []
        for k, v in enumerate(vectors):
            if not isinstance(v, SparseVector):
                raise TypeError(f"Argument {k} was not a sparse vector but was instead {type(v)!r}.")
            idx.append(v._idx + np.uint64(shift))
            val.append(v._val)
            shift += v._n
        if not vectors:
            return cls._make(0, np.zeros(0, np.uint64), np.zeros(0))
        return cls._make(shift, np.concatenate(idx), np.concatenate(val))

    # ---- properties (svector.c:1120-1226) ----
    @property
    def n(self) -> int:
        return self._n

    @n.setter
    def n(self, v) -> None:
        v = int(v)
        if self._idx.size and v <= int(self._idx[-1]):
            raise ValueError(f"Dimension {v} is too small for the entry with the highest index {int(self._idx[-1])}.")
        if v < 0:
            raise ValueError("Dimension can not be negative.")
        self._n = v

    @property
    def values(self) -> np.ndarray:
        return self._val.copy()

    @property
    def indices(self) -> np.ndarray:
        return self._idx.copy()

    @property
    def count(self) -> int:
        return int(self._idx.size)

    @property
    def norm2(self) -> float:
        acc = 0.0
        for v in self._val.tolist():  # same summation order as the C loop
            acc += v * v
        return acc

    # ---- conversions ----
    def __array__(self, dtype=None, copy=None) -> np.ndarray:
        """svec_array, svector.c:435-532: always a fresh dense array."""
        if copy is False:
            raise ValueError("A copy is always created when converting to NDArray.")
        out = np.zeros(self._n, np.float64)
        out[self._idx.astype(np.int64)] = self._val
        return out if dtype is None else out.astype(dtype)

    def __repr__(self) -> str:
        body = ",".join(f" ({int(i)}, {v:g})" for i, v in zip(self._idx, self._val))
        return f"SparseVector({self._n}{',' + body if body else ''})" if body else f"SparseVector({self._n})"

    def __getitem__(self, key):
        """svec_get, svector.c:534-621: unit-step slices give a re-based SparseVector; an integer gives the
        entry or 0.0."""
        if isinstance(key, slice):
            if self._idx.size == 0:
                return self._make(self._n, self._idx, self._val)
            step = 1 if key.step is None else key.step
            if step != 1:
                raise KeyError(f"Sparse vectors do not support slices with steps other than 1 ({step} was given).")
            start, stop, _ = key.indices(self._n)
            length = max(stop - start, 0)
            lo = int(np.searchsorted(self._idx, np.uint64(start), "left"))
            hi = max(int(np.searchsorted(self._idx, np.uint64(max(stop, 0)), "left")), lo)
            return self._make(length, self._idx[lo:hi] - np.uint64(start), self._val[lo:hi])
        i = int(key)
        if i < 0:
            i += self._n
        if i < 0 or i >= self._n:
            raise KeyError(f"Index {int(key)} is outside the allowed range for a vector of dimension {self._n}.")
        pos = int(np.searchsorted(self._idx, np.uint64(i), "left"))
        if pos < self._idx.size and int(self._idx[pos]) == i:
            return float(self._val[pos])
        return 0.0

    # ---- arithmetic (svector.c:1239-1479) ----
    def _check_other(self, other) -> None:
        if not isinstance(other, SparseVector):
            raise TypeError(f"Method requires two sparse vectors, but got a {type(self)!r} and {type(other)!r}.")

    def _combine(self, other: "SparseVector", sign: float) -> "SparseVector":
        if self._n != other._n:
            raise ValueError(f"Can not combine vectors of different dimensions ({self._n} and {other._n}).")
        idx = np.union1d(self._idx, other._idx)
        val = np.zeros(idx.size)
        val[np.searchsorted(idx, self._idx)] = self._val
        pos = np.searchsorted(idx, other._idx)
        val[pos] = val[pos] + sign * other._val
        return self._make(self._n, idx, val)

    def __add__(self, other):
        if not isinstance(other, SparseVector):
            return NotImplemented
        return self._combine(other, 1.0)

    __radd__ = __add__

    def __sub__(self, other):
        if not isinstance(other, SparseVector):
            return NotImplemented
        return self._combine(other, -1.0)

    def __rsub__(self, other):
        if not isinstance(other, SparseVector):
            return NotImplemented
        return other._combine(self, -1.0)

    def __mul__(self, other):
        try:
            k = float(other)
        except (TypeError, ValueError):
            return NotImplemented
        return self._make(self._n, self._idx, self._val * k)

    __rmul__ = __mul__

    def __neg__(self):
        return self._make(self._n, self._idx, -self._val)

    def __eq__(self, other):
        if not isinstance(other, SparseVector):
            return NotImplemented
        return self._n == other._n and np.array_equal(self._idx, other._idx) and np.array_equal(self._val, other._val)

    def __ne__(self, other):
        r = self.__eq__(other)
        return r if r is NotImplemented else not r

    __hash__ = None  # type: ignore[assignment]

    def dot(self, other: "SparseVector") -> float:
        """svec_dot, svector.c:757-818: products of the common entries, summed in index order."""
        self._check_other(other)
        if self._n != other._n:
            raise ValueError(f"Can not compute dot product of vectors of different dimensions ({self._n} and {other._n}).")
        _, ia, ib = np.intersect1d(self._idx, other._idx, assume_unique=True, return_indices=True)
        acc = 0.0
        for a, b in zip(self._val[ia].tolist(), other._val[ib].tolist()):
            acc += a * b
        return acc

    @staticmethod
    def merge_to_dense(*vecs: "SparseVector", duplicates: str = "last") -> np.ndarray:
        """svec_merge_to_dense, svector.c:846-1006: one dense array of the common dimension; entries present
        in several vectors are resolved by ``duplicates`` = "first" | "last" | "sum" | "error"."""
        if not isinstance(duplicates, str):
            raise TypeError(f"Expected keyword value string, got {duplicates!r}.")
        if duplicates not in ("first", "last", "sum", "error"):
            raise ValueError(f"Invalid merge mode '{duplicates}'.")
        if not vecs:
            raise TypeError("Expected at least one sparse vector.")
        for k, v in enumerate(vecs):
            if not isinstance(v, SparseVector):
                raise TypeError(f"Argument {k} was not a sparse vector but was instead {type(v)!r}.")
            if v._n != vecs[0]._n:
                raise ValueError(f"Vector {k} has dimension {v._n}, but the first vector has dimension {vecs[0]._n}.")
        out = np.zeros(vecs[0]._n, np.float64)
        seen = np.zeros(vecs[0]._n, bool)
        for k, v in enumerate(vecs):
            idx = v._idx.astype(np.int64)
            if duplicates == "last":
                out[idx] = v._val
            elif duplicates == "sum":
                out[idx] += v._val
            else:
                dup = seen[idx]
                if duplicates == "error" and dup.any():
                    raise ValueError(f"Duplicate entry at index {int(idx[np.flatnonzero(dup)[0]])} for vector {k}")
                out[idx[~dup]] = v._val[~dup]
            seen[idx] = True
        return out


def _as_u32(a) -> np.ndarray:
    return np.ascontiguousarray(a, np.uint32)


class MatrixCRS:
    """Compressed-row sparse matrix built row by row (crs_matrix.c:38-58, over ``jmtxd_matrix_crs``):
    ``uint32`` column indices sorted within a row, ``float64`` values, row offsets of length ``rows + 1``.
    Everything that reads the matrix requires all rows to have been built (crs_matrix_check_build,
    crs_matrix.c:28-36: ``RuntimeError``)."""

    __slots__ = ("_rows", "_cols", "_built", "_off", "_idx", "_val")

    def __init__(self, rows: int, cols: int):
        rows, cols = int(rows), int(cols)
        if rows < 0 or cols < 0:
            raise OverflowError("can't convert negative value to unsigned int")
        self._rows, self._cols, self._built = rows, cols, 0
        self._off = np.zeros(rows + 1, np.int64)
        self._idx = np.zeros(0, np.uint32)
        self._val = np.zeros(0, np.float64)

    @classmethod
    def _from_csr(cls, rows: int, cols: int, off, idx, val) -> "MatrixCRS":
        self = cls(rows, cols)
        self._off = np.ascontiguousarray(off, np.int64)
        self._idx, self._val = _as_u32(idx), np.ascontiguousarray(val, np.float64)
        self._built = rows
        return self

    def _check_build(self) -> None:
        if self._built != self._rows:
            raise RuntimeError(f"Matrix has only {self._built} of {self._rows} rows built.")

    # ---- construction ----
    def build_row(self, row: int, entries: SparseVector | None = None) -> None:
        """crs_matrix_build_row, crs_matrix.c:68-152: rows are set in order; building row ``r`` again drops
        every row after it."""
        row = int(row)
        if row < 0:
            raise OverflowError("can't convert negative value to unsigned int")
        if row >= self._rows:
            raise ValueError(f"Row index {row} out of bounds for matrix of dimensions ({self._rows}, {self._cols})")
        if row > self._built:
            raise ValueError(f"Row index {row} is greater than the number of rows built ({self._built})")
        if entries is not None and not isinstance(entries, SparseVector):
            raise TypeError(f"entries must be a SparseVector, got {type(entries)!r}")
        self._built = row + 1
        begin = int(self._off[row])
        if entries is None:
            idx, val = np.zeros(0, np.uint32), np.zeros(0)
        else:
            if entries._n != self._cols:
                raise ValueError(f"Dimension of the sparse vector ({entries._n}) does not match number of columns ({self._cols})")
            idx, val = entries._idx.astype(np.uint32), entries._val
        self._idx = np.concatenate([self._idx[:begin], idx])
        self._val = np.concatenate([self._val[:begin], val])
        self._off[row + 1:] = begin + idx.size

    def set_from_data(self, values, column_indices, row_lengths) -> None:
        """crs_matrix_from_data, crs_matrix.c:689-804 (the scipy ``csr_array`` exchange format)."""
        val = np.asarray(values, np.float64)
        idx = np.asarray(column_indices)
        lens = np.asarray(row_lengths)
        if val.ndim != 1 or idx.ndim != 1 or lens.ndim != 1:
            raise ValueError("values, column_indices and row_lengths must be one dimensional")
        if idx.size != val.size:
            raise ValueError(f"values and column_indices must have the same length ({val.size} and {idx.size}).")
        if lens.size != self._rows:
            raise ValueError(f"row_lengths must have one entry per row ({self._rows}), got {lens.size}.")
        if lens.size and np.any(lens < 0):
            raise ValueError("row lengths can not be negative")
        off = np.concatenate([[0], np.cumsum(lens.astype(np.int64))])
        if off[-1] != val.size:
            raise ValueError(f"Sum of row lengths ({int(off[-1])}) does not match the number of values ({val.size}).")
        if idx.size:
            if np.any(idx < 0) or np.any(idx >= self._cols):
                raise ValueError(f"Column indices must be in the range [0, {self._cols}).")
            inner = np.ones(idx.size, bool)
            inner[off[:-1][off[:-1] < idx.size]] = False
            if np.any((idx[1:] <= idx[:-1]) & inner[1:]):
                raise ValueError("Column indices must be sorted within each row.")
        self._off, self._idx, self._val, self._built = off, _as_u32(idx).copy(), val.copy(), self._rows

    @classmethod
    def from_dense(cls, x) -> "MatrixCRS":
        """crs_matrix_from_dense, crs_matrix.c:1029-1070: every entry is stored, zeros included."""
        a = np.asarray(x, np.float64)
        if a.ndim != 2:
            raise ValueError(f"Expected a two dimensional array, got {a.ndim} dimensions.")
        r, c = a.shape
        return cls._from_csr(r, c, np.arange(r + 1, dtype=np.int64) * c, np.tile(np.arange(c, dtype=np.uint32), r), a.ravel().copy())

    # ---- readers ----
    @property
    def shape(self) -> tuple[int, int]:
        return (self._rows, self._cols)

    @property
    def built_rows(self) -> int:
        return self._built

    @property
    def column_indices(self) -> np.ndarray:
        self._check_build()
        return self._idx.copy()

    @property
    def row_indices(self) -> np.ndarray:
        self._check_build()
        return np.repeat(np.arange(self._rows, dtype=np.uint32), np.diff(self._off))

    @property
    def values(self) -> np.ndarray:
        self._check_build()
        return self._val.copy()

    @property
    def row_offsets(self) -> np.ndarray:
        self._check_build()
        return self._off.astype(np.uint32)

    @property
    def position_pairs(self) -> np.ndarray:
        self._check_build()
        return np.stack([self.row_indices, self._idx], axis=1).astype(np.uint32).reshape(-1, 2)

    @property
    def nonempty_rows(self) -> np.ndarray:
        self._check_build()
        return np.flatnonzero(np.diff(self._off) > 0).astype(np.uintc)

    def toarray(self) -> np.ndarray:
        self._check_build()
        out = np.zeros((self._rows, self._cols), np.float64)
        out[self.row_indices.astype(np.int64), self._idx.astype(np.int64)] = self._val
        return out

    def __array__(self, dtype=None, copy=None) -> np.ndarray:
        out = self.toarray()
        return out if dtype is None else out.astype(dtype)

    def _row(self, r: int):
        return self._idx[self._off[r]:self._off[r + 1]], self._val[self._off[r]:self._off[r + 1]]

    def __getitem__(self, key):
        """crs_matrix_get_row, crs_matrix.c:1371-1426: ``m[r]`` is the row as a SparseVector, ``m[r, c]`` one
        entry."""
        self._check_build()
        if isinstance(key, tuple):
            if len(key) != 2:
                raise KeyError(f"Expected a row index or a (row, column) pair, got {key!r}")
            r, c = int(key[0]), int(key[1])
            if not (0 <= r < self._rows and 0 <= c < self._cols):
                raise KeyError(f"Index ({r}, {c}) is outside the matrix of shape {self.shape}.")
            idx, val = self._row(r)
            pos = int(np.searchsorted(idx, c))
            return float(val[pos]) if pos < idx.size and int(idx[pos]) == c else 0.0
        r = int(key)
        if not 0 <= r < self._rows:
            raise KeyError(f"Row index {r} is outside the allowed range [0, {self._rows}).")
        idx, val = self._row(r)
        return SparseVector._make(self._cols, idx.astype(np.uint64), val.copy())

    def __str__(self) -> str:
        return f"MatrixCRS({self._rows}, {self._cols})"

    __repr__ = __str__

    # ---- in-place edits ----
    def shrink(self) -> None:
        self._check_build()  # storage is always exact here

    def remove_entries_bellow(self, v: float = 0.0) -> int:
        """crs_matrix_remove_entries_bellow, crs_matrix.c:948-984: drops ``|a_ij| < v``, returns the count."""
        v = float(v)
        self._check_build()
        if v < 0:
            raise ValueError(f"Value must be non-negative, got {v:f}.")
        keep = ~(np.abs(self._val) < v)
        removed = int(keep.size - keep.sum())
        rows = self.row_indices.astype(np.int64)[keep]
        self._idx, self._val = self._idx[keep], self._val[keep]
        self._off = np.concatenate([[0], np.cumsum(np.bincount(rows, minlength=self._rows))]).astype(np.int64)
        return removed

    def add_to_dense(self, out: np.ndarray) -> None:
        """crs_matrix_add_to_dense, crs_matrix.c:986-1027: ``out`` must be a writable C-contiguous float64
        array of exactly the matrix's shape."""
        self._check_build()
        if not isinstance(out, np.ndarray) or out.dtype != np.float64:
            raise ValueError("out must be a float64 array")
        if out.shape != self.shape:
            raise ValueError(f"out must have the shape {self.shape}, got {out.shape}")
        if not (out.flags.c_contiguous and out.flags.writeable):
            raise ValueError("out must be C-contiguous and writable")
        np.add.at(out, (self.row_indices.astype(np.int64), self._idx.astype(np.int64)), self._val)

    def transpose(self) -> "MatrixCRS":
        self._check_build()
        rows = self.row_indices.astype(np.int64)
        order = np.lexsort((rows, self._idx))
        off = np.concatenate([[0], np.cumsum(np.bincount(self._idx.astype(np.int64), minlength=self._cols))])
        return self._from_csr(self._cols, self._rows, off, rows[order], self._val[order])

    # ---- products ----
    def multiply_to_sparse(self, x) -> SparseVector:
        """crs_matrix_multiply_to_sparse, crs_matrix.c:1086-1160: ``A x`` keeping only non-zero results."""
        self._check_build()
        x = np.asarray(x, np.float64)
        if x.ndim != 1 or x.size != self._cols:
            raise ValueError(f"Expected a one dimensional array with {self._cols} entries, got shape {x.shape}.")
        idx, val = [], []
        for r in np.flatnonzero(np.diff(self._off) > 0):
            ci, cv = self._row(int(r))
            acc = 0.0
            for a, b in zip(cv.tolist(), x[ci.astype(np.int64)].tolist()):
                acc += a * b
            if acc != 0.0:
                idx.append(int(r))
                val.append(acc)
        return SparseVector._make(self._rows, np.array(idx, np.uint64), np.array(val, np.float64))

    def _matvec_sparse(self, v: SparseVector) -> SparseVector:
        # crs_matrix_matmul_left_svec, crs_matrix.c:205-270
        if self._cols != v._n:
            raise ValueError("Dimension of the sparse vector does not match number of columns")
        dense = np.zeros(self._cols)
        mask = np.zeros(self._cols, bool)
        dense[v._idx.astype(np.int64)] = v._val
        mask[v._idx.astype(np.int64)] = True
        idx, val = [], []
        for r in np.flatnonzero(np.diff(self._off) > 0):
            ci, cv = self._row(int(r))
            ci = ci.astype(np.int64)
            acc = 0.0
            for a, b, m in zip(cv.tolist(), dense[ci].tolist(), mask[ci].tolist()):
                if m:
                    acc += a * b
            if acc != 0.0:
                idx.append(int(r))
                val.append(acc)
        return SparseVector._make(self._rows, np.array(idx, np.uint64), np.array(val, np.float64))

    def _rmatvec_sparse(self, v: SparseVector) -> SparseVector:
        # crs_matrix_matmul_right_svec, crs_matrix.c:272-338: rows scaled by the vector's entries, added up in
        # the order of the entries; exact-zero products are skipped
        if self._rows != v._n:
            raise ValueError("Dimension of the sparse vector does not match number of rows")
        acc = np.zeros(self._cols)
        hit = np.zeros(self._cols, bool)
        for i, e in zip(v._idx.tolist(), v._val.tolist()):
            if e == 0.0:
                continue
            ci, cv = self._row(int(i))
            prod = cv * e
            nz = prod != 0.0
            ci = ci[nz].astype(np.int64)
            acc[ci] += prod[nz]
            hit[ci] = True
        idx = np.flatnonzero(hit)
        return SparseVector._make(self._cols, idx.astype(np.uint64), acc[idx])

    def _matmat(self, other: "MatrixCRS") -> "MatrixCRS":
        # crs_matrix_matmul_double, crs_matrix.c:340-494: row r of the product = sum_j a_rj * (row j of other)
        if self._cols != other._rows:
            raise ValueError(
                f"Can not multiply matrices unless the number of columns of the left matrix ({self._cols}) is equal to "
                f"the number of rows of the right matrix ({other._rows})."
            )
        off, idx, val = [0], [], []
        for r in range(self._rows):
            ci, cv = self._row(r)
            row = other._rmatvec_sparse(SparseVector._make(other._rows, ci.astype(np.uint64), cv))
            idx.append(row._idx.astype(np.uint32))
            val.append(row._val)
            off.append(off[-1] + row._idx.size)
        return self._from_csr(self._rows, other._cols, np.array(off, np.int64),
                              np.concatenate(idx) if idx else np.zeros(0, np.uint32), np.concatenate(val) if val else np.zeros(0))

    def __matmul__(self, other):
        """crs_matrix_matmul / _matmul_left, crs_matrix.c:496-673: CRS @ {MatrixCRS, SparseVector, ndarray}."""
        self._check_build()
        if isinstance(other, MatrixCRS):
            other._check_build()
            return self._matmat(other)
        if isinstance(other, SparseVector):
            return self._matvec_sparse(other)
        x = np.asarray(other, np.float64)
        if x.ndim not in (1, 2) or x.shape[0] != self._cols:
            raise ValueError(f"Can not multiply a matrix of shape {self.shape} with an array of shape {x.shape}.")
        out = np.zeros((self._rows,) + x.shape[1:])
        np.add.at(out, self.row_indices.astype(np.int64), self._val.reshape((-1,) + (1,) * (x.ndim - 1)) * x[self._idx.astype(np.int64)])
        return out

    def __rmatmul__(self, other):
        """crs_matrix_matmul_right, crs_matrix.c:569-647: {SparseVector, ndarray} @ CRS."""
        self._check_build()
        if isinstance(other, SparseVector):
            return self._rmatvec_sparse(other)
        x = np.asarray(other, np.float64)
        if x.ndim not in (1, 2) or x.shape[-1] != self._rows:
            raise ValueError(f"Can not multiply an array of shape {x.shape} with a matrix of shape {self.shape}.")
        out = np.zeros(x.shape[:-1] + (self._cols,))
        rows = self.row_indices.astype(np.int64)
        np.add.at(out.T, self._idx.astype(np.int64), (x[..., rows] * self._val).T)
        return out

    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        """crs_matrix_array_ufunc, crs_matrix.c:806-869: only ``ndarray @ MatrixCRS`` / ``MatrixCRS @ ndarray``
        are taken over from numpy."""
        if ufunc is np.matmul and method == "__call__" and len(inputs) == 2 and not kwargs:
            a, b = inputs
            if a is self:
                return self.__matmul__(b)
            if b is self:
                return self.__rmatmul__(a)
        return NotImplemented

    # ---- flat triples for the device library ----
    def coo(self) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
        """``(row, column, value)`` of all stored entries in row-major order (not part of the reference type)."""
        self._check_build()
        return self.row_indices.astype(np.int64), self._idx.astype(np.int64), self._val.copy()


def constraint_blocks_to_coo(blocks: Iterable[MatrixCRS], offsets: np.ndarray):
    """Per-element constraint blocks ``N_e`` (``n_trace x N_e`` MatrixCRS, python/mfv2d/solving.py:62-90) ->
    global COO over the packed dense columns, sorted by (row, column)."""
    rows, cols, vals = [], [], []
    for e, m in enumerate(blocks):
        r, c, v = m.coo()
        rows.append(r)
        cols.append(c + int(offsets[e]))
        vals.append(v)
    if not rows:
        return np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0)
    row, col, val = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
    order = np.lexsort((col, row))
    return row[order], col[order], val[order]
