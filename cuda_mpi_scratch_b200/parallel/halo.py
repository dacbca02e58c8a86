"""Layout descriptors, tile regions and the unfused 8-neighbour halo exchange.

Python face of ``include/b200comm/stencil2d.hpp`` -- the capability of the
reference's ``stencil2d/stencil2D.h`` (``Array2D`` :30-42, ``RegionID`` :79-82,
``SubArrayRegion`` :107-201, ``CreateSendRecvArrays`` :381-437, ``ExchangeData``
:363-377, ``Print`` :92-102, ``PrintCartesianGrid`` :513-530).  The exchange
uses subarray datatypes over the comm layer, so it works on host tensors (CPU
tests, no GPU) and on CUDA tensors (moved by the pack/move kernel).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

from .. import _lib
from .comm import Comm, _bufinfo
from .datatype import Datatype

_L = _lib.lib

(TOP_LEFT, TOP_CENTER, TOP_RIGHT, CENTER_LEFT, CENTER, CENTER_RIGHT, BOTTOM_LEFT, BOTTOM_CENTER,
 BOTTOM_RIGHT, TOP, LEFT, BOTTOM, RIGHT) = range(13)
REGION_NAMES = ("top left", "top center", "top right", "center left", "center", "center right",
                "bottom left", "bottom center", "bottom right", "top", "left", "bottom", "right")


class CArray2D(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("x_offset", C.c_int),
                ("y_offset", C.c_int), ("row_stride", C.c_int)]


_lib._proto("b2c_sub_array_region", [C.POINTER(CArray2D), C.c_int, C.c_int, C.c_int,
                                     C.POINTER(CArray2D)])
_lib._proto("b2c_halo_plan_create", [C.c_void_p, C.POINTER(CArray2D), C.c_int, C.c_int, C.c_void_p,
                                     C.POINTER(C.c_void_p)])
_lib._proto("b2c_halo_plan_destroy", [C.c_void_p])
_lib._proto("b2c_halo_exchange", [C.c_void_p, C.c_void_p])
_lib._proto("b2c_halo_plan_entry", [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int),
                                    C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(CArray2D),
                                    C.POINTER(C.c_int64)])
_lib._proto("b2c_region_report", [C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int])


@dataclass(frozen=True)
class Array2D:
    """Pure layout descriptor (the data pointer is separate): ``width`` x ``height``
    elements at ``(x_offset, y_offset)`` of an allocation with ``row_stride``
    elements per row.  x is the fast axis."""
    width: int
    height: int
    row_stride: int
    x_offset: int = 0
    y_offset: int = 0

    def _c(self) -> CArray2D:
        return CArray2D(self.width, self.height, self.x_offset, self.y_offset, self.row_stride)

    @staticmethod
    def _from(c: CArray2D) -> "Array2D":
        return Array2D(c.width, c.height, c.row_stride, c.x_offset, c.y_offset)

    def view(self, tensor):
        """The region as a view of a 2-D tensor that covers the allocation."""
        return tensor[self.y_offset:self.y_offset + self.height,
                      self.x_offset:self.x_offset + self.width]


def sub_array_region(g: Array2D, stencil_w: int, stencil_h: int, region: int) -> Array2D:
    out = CArray2D()
    _lib.check(_L.b2c_sub_array_region(C.byref(g._c()), stencil_w, stencil_h, region, C.byref(out)))
    return Array2D._from(out)


def region_report(w: int, h: int, stencil_w: int, stencil_h: int) -> str:
    """Text listing of every region (the reference's ``TestSubRegionExtraction``)."""
    b = C.create_string_buffer(8192)
    _lib.check(_L.b2c_region_report(w, h, stencil_w, stencil_h, b, 8192))
    return b.value.decode()


class HaloPlan:
    """8 sends + 8 receives of subarray types on a 2-D Cartesian communicator."""

    def __init__(self, cart: Comm, tile: Array2D, stencil_w: int, stencil_h: int, elem: Datatype):
        self.cart, self.tile, self.elem = cart, tile, elem
        self.stencil = (stencil_w, stencil_h)
        h = C.c_void_p()
        _lib.check(_L.b2c_halo_plan_create(cart._h, C.byref(tile._c()), stencil_w, stencil_h,
                                           elem._h, C.byref(h)))
        self._h = h

    @property
    def core(self) -> Array2D:
        return sub_array_region(self.tile, *self.stencil, CENTER)

    def entries(self, which: str) -> list[dict]:
        out = []
        for i in range(8):
            peer, tag, reg, nb = C.c_int(), C.c_int(), C.c_int(), C.c_int64()
            lay = CArray2D()
            _lib.check(_L.b2c_halo_plan_entry(self._h, 1 if which == "send" else 0, i, C.byref(peer),
                                              C.byref(tag), C.byref(reg), C.byref(lay), C.byref(nb)))
            out.append({"peer": peer.value, "tag": tag.value, "region": reg.value,
                        "layout": Array2D._from(lay), "bytes": nb.value})
        return out

    def exchange(self, tile_tensor) -> None:
        """All receives, all sends, one waitall (``ExchangeData``)."""
        addr, _, _, keep = _bufinfo(tile_tensor)
        _lib.check(_L.b2c_halo_exchange(self._h, addr))

    def close(self) -> None:
        if self._h:
            _L.b2c_halo_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


def print_array(t, fmt: str = "{:g}") -> str:
    """Row-major dump, every value followed by a blank (``Print``)."""
    return "".join("".join(fmt.format(v) + " " for v in row) + "\n" for row in t.tolist())


def print_cart_grid(cart: Comm) -> str:
    dims, _ = cart.cart_get()
    return "".join("".join(f"{cart.cart_rank([r, c])} " for c in range(dims[1])) + "\n"
                   for r in range(dims[0]))
