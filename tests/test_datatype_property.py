"""Property test of the datatype engine against a pure-Python typemap model.

The model follows the MPI definitions directly: a datatype is a sequence of (byte displacement,
byte length) entries in typemap order plus a lower bound and an extent; constructors compose
them.  For random nested types (up to three levels of contiguous / vector / hvector / indexed /
hindexed / subarray / struct / resized) the native engine must agree on size, bounds, the exact
byte sequence of ``count`` elements, and pack/unpack must gather/scatter exactly those bytes.
The reference relies on an MPI library for all of this (mpi7.cpp:32-41, mpi8.cpp:38-54,
mpi-complex-types.cpp:33-50, stencil2d/stencil2D.h:210-228).
"""
from __future__ import annotations

import itertools

import numpy as np
from hypothesis import HealthCheck, given, settings, strategies as st

from cuda_mpi_scratch_b200.parallel.datatype import ORDER_C, ORDER_FORTRAN, Datatype

BASICS = {"int8": 1, "int16": 2, "int32": 4, "float64": 8}


class Model:
    """(entries in typemap order, lb, extent) + the native handle built the same way."""

    def __init__(self, entries, lb, extent, native):
        self.entries, self.lb, self.extent, self.native = entries, lb, extent, native

    @property
    def size(self):
        return sum(n for _, n in self.entries)

    def bytes_of(self, count=1):
        out = []
        for i in range(count):
            for d, n in self.entries:
                out.extend(range(d + i * self.extent, d + i * self.extent + n))
        return out


def basic(kind):
    s = BASICS[kind]
    return Model([(0, s)], 0, s, Datatype.basic(kind))


def _bounds(placed, old):
    """lb/ub of copies of `old` placed at the byte displacements `placed`."""
    lb = min(d + old.lb for d in placed)
    ub = max(d + old.lb + old.extent for d in placed)
    return lb, ub - lb


def _place(old, disps):
    entries = [(d + e, n) for d in disps for e, n in old.entries]
    return entries


def contiguous(old, count):
    disps = [i * old.extent for i in range(count)]
    lb, ext = _bounds(disps, old)
    return Model(_place(old, disps), lb, ext, old.native.contiguous(count))


def vector(old, count, blocklen, stride):
    disps = [(i * stride + j) * old.extent for i in range(count) for j in range(blocklen)]
    lb, ext = _bounds(disps, old)
    return Model(_place(old, disps), lb, ext, old.native.vector(count, blocklen, stride))


def hvector(old, count, blocklen, stride_bytes):
    disps = [i * stride_bytes + j * old.extent for i in range(count) for j in range(blocklen)]
    lb, ext = _bounds(disps, old)
    return Model(_place(old, disps), lb, ext, old.native.hvector(count, blocklen, stride_bytes))


def indexed(old, blocklens, disps_el):
    disps = [(d + j) * old.extent for b, d in zip(blocklens, disps_el) for j in range(b)]
    lb, ext = _bounds(disps, old)
    return Model(_place(old, disps), lb, ext, old.native.indexed(blocklens, disps_el))


def hindexed(old, blocklens, disps_b):
    disps = [d + j * old.extent for b, d in zip(blocklens, disps_b) for j in range(b)]
    lb, ext = _bounds(disps, old)
    return Model(_place(old, disps), lb, ext, old.native.hindexed(blocklens, disps_b))


def struct(olds, blocklens, disps_b):
    entries, lbs, ubs = [], [], []
    for old, b, d in zip(olds, blocklens, disps_b):
        ds = [d + j * old.extent for j in range(b)]
        entries += _place(old, ds)
        lo, ext = _bounds(ds, old)
        lbs.append(lo)
        ubs.append(lo + ext)
    return Model(entries, min(lbs), max(ubs) - min(lbs), Datatype.struct(blocklens, disps_b, [o.native for o in olds]))


def subarray(old, sizes, subsizes, starts, order):
    nd = len(sizes)
    strides = [0] * nd                      # element strides of the full array
    if order == ORDER_C:
        acc = 1
        for k in reversed(range(nd)):
            strides[k] = acc
            acc *= sizes[k]
        idx_iter = itertools.product(*[range(s) for s in subsizes])
        disps = [sum((starts[k] + ix[k]) * strides[k] for k in range(nd)) * old.extent for ix in idx_iter]
    else:
        acc = 1
        for k in range(nd):
            strides[k] = acc
            acc *= sizes[k]
        # typemap order: first index fastest
        idx_iter = itertools.product(*[range(s) for s in reversed(subsizes)])
        disps = [sum((starts[k] + ix[nd - 1 - k]) * strides[k] for k in range(nd)) * old.extent for ix in idx_iter]
    total = int(np.prod(sizes))
    return Model(_place(old, disps), 0, total * old.extent, old.native.subarray(sizes, subsizes, starts, order))


def resized(old, lb, extent):
    return Model(list(old.entries), lb, extent, old.native.resized(lb, extent))


@st.composite
def derived(draw, depth=0):
    if depth >= 3 or draw(st.integers(0, 4)) == 0:
        return basic(draw(st.sampled_from(sorted(BASICS))))
    old = draw(derived(depth=depth + 1))
    if old.extent <= 0 or old.size == 0 or old.size > 600:
        return old
    kind = draw(st.sampled_from(["contiguous", "vector", "hvector", "indexed", "hindexed", "subarray",
                                 "struct", "resized"]))
    if kind == "contiguous":
        return contiguous(old, draw(st.integers(1, 4)))
    if kind == "vector":
        bl = draw(st.integers(1, 3))
        return vector(old, draw(st.integers(1, 4)), bl, draw(st.integers(bl, bl + 3)))
    if kind == "hvector":
        bl = draw(st.integers(1, 3))
        return hvector(old, draw(st.integers(1, 4)), bl, bl * old.extent + draw(st.integers(0, 24)))
    if kind == "indexed":
        n = draw(st.integers(1, 3))
        bls = [draw(st.integers(1, 3)) for _ in range(n)]
        ds, pos = [], draw(st.integers(0, 3))
        for b in bls:                      # increasing, non-overlapping blocks
            ds.append(pos)
            pos += b + draw(st.integers(0, 3))
        return indexed(old, bls, ds)
    if kind == "hindexed":
        n = draw(st.integers(1, 3))
        bls = [draw(st.integers(1, 2)) for _ in range(n)]
        ds, pos = [], draw(st.integers(-64, 64))     # signed byte displacements
        for b in bls:
            ds.append(pos)
            pos += b * old.extent + draw(st.integers(0, 40))
        if draw(st.booleans()):
            ds, bls = ds[::-1], bls[::-1]            # typemap order need not be address order
        return hindexed(old, bls, ds)
    if kind == "subarray":
        nd = draw(st.integers(1, 3))
        sizes = [draw(st.integers(1, 5)) for _ in range(nd)]
        subs = [draw(st.integers(1, s)) for s in sizes]
        starts = [draw(st.integers(0, s - u)) for s, u in zip(sizes, subs)]
        return subarray(old, sizes, subs, starts, draw(st.sampled_from([ORDER_C, ORDER_FORTRAN])))
    if kind == "struct":
        other = basic(draw(st.sampled_from(sorted(BASICS))))
        b0, b1 = draw(st.integers(1, 2)), draw(st.integers(1, 3))
        d1 = b0 * old.extent + max(0, old.lb) + draw(st.integers(0, 16))
        return struct([old, other], [b0, b1], [0, d1])
    span = max(d + n for d, n in old.entries) - min(d for d, _ in old.entries)
    return resized(old, draw(st.integers(-8, 8)), span + draw(st.integers(0, 32)))


def _expand(runs):
    out = []
    for d, n in runs:
        out.extend(range(d, d + n))
    return out


@settings(max_examples=200, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(derived(), st.integers(1, 3))
def test_native_engine_agrees_with_the_typemap_model(m, count):
    t = m.native.commit()
    assert t.size == m.size
    assert (t.lb, t.extent) == (m.lb, m.extent), t.describe()
    want = m.bytes_of(count)
    assert _expand(t.runs(count)) == want, t.describe()
    # runs are maximal: two consecutive runs never touch
    runs = t.runs(count)
    assert all(a[0] + a[1] != b[0] for a, b in zip(runs, runs[1:]))
    # pack gathers exactly those bytes in typemap order; unpack scatters them back
    if len(set(want)) != len(want):
        return                                  # overlapping typemaps are legal to send, not to receive
    lo, hi = min(want), max(want) + 1
    rng = np.random.default_rng(len(want) * 31 + count)
    buf = rng.integers(0, 256, hi - lo + 16, dtype=np.uint8)
    base = buf.ctypes.data - lo + 8             # buffer address such that displacement `lo` is buf[8]
    packed = t.pack(base, count)
    assert list(packed) == [int(buf[8 + o - lo]) for o in want]
    out = np.zeros_like(buf)
    t.unpack(bytes(packed), out.ctypes.data - lo + 8, count)
    expect = np.zeros_like(buf)
    for o in want:
        expect[8 + o - lo] = buf[8 + o - lo]
    assert np.array_equal(out, expect)
