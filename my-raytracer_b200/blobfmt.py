"""ctypes mirror of the flattened blob's header (csrc/rt_types.h BlobHeader): lets the host side inspect a blob it
is about to broadcast or has received (section offsets and record counts) without the library."""
from __future__ import annotations

import ctypes as C

BLOB_MAGIC = 0x42325452
BLOB_VERSION = 12


class BlobHeader(C.Structure):
    _fields_ = [
        ("magic", C.c_uint32), ("version", C.c_uint32), ("total_bytes", C.c_uint64),
        ("num_nodes", C.c_int32), ("num_tris", C.c_int32), ("num_meshes", C.c_int32), ("num_objects", C.c_int32),
        ("root_is_leaf", C.c_int32), ("max_depth", C.c_int32), ("max_leaf", C.c_int32), ("num_nodes_ref", C.c_int32),
        ("num_texels", C.c_uint64),
        ("off_nodes", C.c_uint64), ("off_filter", C.c_uint64), ("off_exact", C.c_uint64), ("off_shade", C.c_uint64),
        ("off_meshes", C.c_uint64), ("off_objects", C.c_uint64), ("off_texels", C.c_uint64),
        ("scene_extent", C.c_float), ("box_extent", C.c_float), ("pad_f", C.c_float * 2),
        ("off_maxx", C.c_uint64), ("root_max_x", C.c_double),
        ("off_tri_id", C.c_uint64), ("off_tri_pos", C.c_uint64),
        ("q_base", C.c_int32), ("num_nodes_q", C.c_int32), ("topology", C.c_int32), ("num_nodes4", C.c_int32),
        ("off_objfilt", C.c_uint64), ("off_obj_order", C.c_uint64), ("off_lin_idx", C.c_uint64),
        ("obj_root", C.c_int32), ("num_nodes_obj", C.c_int32), ("num_obj_tree", C.c_int32), ("num_lin", C.c_int32),
        ("off_nodes4", C.c_uint64),
        ("num_nodes4_tri", C.c_int32), ("obj_root4", C.c_int32), ("pad_h", C.c_int32 * 2),
        ("off_nodes8", C.c_uint64), ("num_nodes8", C.c_int32), ("pad_8", C.c_int32),
    ]


def header(blob) -> dict:
    """Header fields of a blob (bytes-like) as a dict."""
    h = BlobHeader.from_buffer_copy(bytes(blob[:C.sizeof(BlobHeader)]))
    return {n: getattr(h, n) for n, _ in BlobHeader._fields_ if not n.startswith("pad")}


def poke_u64(blob: bytearray, field: str, value: int) -> None:
    """Overwrite a 64-bit header field in place (tests: corrupt a section offset)."""
    off = getattr(BlobHeader, field).offset
    blob[off:off + 8] = int(value).to_bytes(8, "little")
