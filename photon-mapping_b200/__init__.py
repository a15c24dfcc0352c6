"""photon-mapping_b200 — Python host binding of libpmhot (the B200-native photon-mapping hot path).

The directory name carries a hyphen (it is the repository's mandated layout), so import it through
``__graft_entry__.load_package()`` which registers it as ``photon_mapping_b200``.

This module is plumbing only: ctypes declarations for every symbol of ``include/pmhot.h`` plus thin
wrappers that accept numpy arrays (host memory) or torch CUDA tensors (device memory) and hand raw
pointers to the C ABI.  There is NO CPU fallback and nothing here touches the test oracle: if the CUDA
library is missing or no GPU is usable the calls raise.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_uint8, c_uint32, c_uint64, c_void_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpmhot.so")

# ---- record layouts (byte-for-byte include/pmhot.h) -------------------------------------------
PHOTON40 = np.dtype([("pos", "<f4", 3), ("dir", "<f4", 3), ("power", "<i4"), ("color", "<f4", 3)])
PHOTON44 = np.dtype(
    [("pos", "<f4", 3), ("dir", "<f4", 3), ("color", "<f4", 3), ("power", "<f4"), ("qn", "u1", 3), ("split_dim", "u1")]
)
MATERIAL = np.dtype(
    [("albedo", "<f4", 3), ("diffuse", "<f4"), ("specular", "<f4"), ("transmission", "<f4"), ("refraction_idx", "<f4")]
)
LIGHT = np.dtype([("pos", "<f4", 3), ("rgb", "<f4", 3), ("power", "<f8")])
LIGHT_SHAPE = np.dtype([("type", "<i4"), ("normal", "<f4", 3), ("side_length", "<f4")])
LIGHT_POINT, LIGHT_SQUARE = 0, 1
CAMERA = np.dtype([("pos", "<f4", 3), ("dir_00", "<f4", 3), ("dir_du", "<f4", 3), ("dir_dv", "<f4", 3)])
RAY = np.dtype([("org", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4")])
HIT = np.dtype([("prim", "<i4"), ("t", "<f4")])
assert PHOTON40.itemsize == 40 and PHOTON44.itemsize == 44 and MATERIAL.itemsize == 28
assert LIGHT.itemsize == 32 and CAMERA.itemsize == 48 and RAY.itemsize == 32 and HIT.itemsize == 8

RNG_LCG_PARITY = 0
RNG_COUNTER = 1

SUPPORTED_K = (8, 16, 32, 50, 64)
KNN_STATIC_MAX_QUERIES = 262144  # PM_KNN_STATIC_MAX_QUERIES

# every exported symbol of include/pmhot.h: name -> (restype, argtypes)
_SIGNATURES = {
    "pm_last_error": (c_char_p, []),
    "pm_version": (c_int, []),
    "pm_ctx_create": (c_int, [c_int, POINTER(c_void_p)]),
    "pm_ctx_destroy": (None, [c_void_p]),
    "pm_ctx_stream": (c_void_p, [c_void_p]),
    "pm_ctx_sync": (c_int, [c_void_p]),
    "pm_ctx_wait_stream": (c_int, [c_void_p, c_void_p]),
    "pm_ctx_launch_count": (c_int64, [c_void_p]),
    "pm_ctx_reset_launch_count": (None, [c_void_p]),
    "pm_ctx_last_kernel_ms": (c_float, [c_void_p]),
    "pm_ctx_last_fallback_count": (c_int64, [c_void_p]),
    "pm_ctx_last_handover_count": (c_int64, [c_void_p]),
    "pm_scene_create": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, POINTER(c_void_p)]),
    "pm_scene_destroy": (None, [c_void_p]),
    "pm_trace_closest": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "pm_trace_occluded": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "pm_emit_photons": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int, c_uint64, c_int, c_int64, c_int64, c_void_p,
         c_int64, POINTER(c_int64), POINTER(c_int64), POINTER(c_int64)],
    ),
    "pm_emit_photons_keyed": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int, c_uint64, c_int, c_int64, c_int64, c_void_p,
         c_void_p, c_int64, POINTER(c_int64), POINTER(c_int64), POINTER(c_int64)],
    ),
    "pm_emit_photons_shaped": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int, c_uint64, c_int, c_int64, c_int64, c_void_p,
         c_void_p, c_int64, POINTER(c_int64), POINTER(c_int64), POINTER(c_int64)],
    ),
    "pm_emit_photons_counted": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int, c_uint64, c_int, c_int64, c_int64, c_void_p,
         c_void_p, c_int64, POINTER(c_int64), POINTER(c_int64), POINTER(c_int64)],
    ),
    "pm_dev_alloc": (c_int, [c_void_p, c_int64, POINTER(c_void_p)]),
    "pm_dev_free": (None, [c_void_p, c_void_p]),
    "pm_dev_copy": (c_int, [c_void_p, c_void_p, c_void_p, c_int64]),
    "pm_device_count": (c_int, []),
    "pm_comm_unique_id": (c_int, [c_void_p, c_int]),
    "pm_comm_create": (c_int, [c_void_p, c_void_p, c_int, c_int, POINTER(c_void_p)]),
    "pm_comm_destroy": (None, [c_void_p]),
    "pm_comm_rank": (c_int, [c_void_p]),
    "pm_comm_world": (c_int, [c_void_p]),
    "pm_comm_allreduce_sum": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int]),
    "pm_comm_allgather": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64]),
    "pm_exchange_photons": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p, c_int64, c_void_p, c_float,
                                    c_void_p, c_int64, POINTER(c_int64), POINTER(c_float)]),
    "pm_exchange_place": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_float,
                                  c_void_p, c_int64, POINTER(c_int64)]),
    "pm_map_build_sharded": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, POINTER(c_void_p)]),
    "pm_ppm_update": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                              POINTER(c_float)]),
    "pm_photons_to_maps": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p]),
    "pm_map_build": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, POINTER(c_void_p)]),
    "pm_test_build_carry_from": (None, [c_int64]),
    "pm_map_from_tree": (c_int, [c_void_p, c_void_p, c_int64, POINTER(c_void_p)]),
    "pm_map_build_shard": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "pm_map_build_finish": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, POINTER(c_void_p)]),
    "pm_map_build_cells": (c_int, [c_void_p, c_void_p, c_int64, c_float, POINTER(c_void_p)]),
    "pm_map_destroy": (None, [c_void_p]),
    "pm_map_size": (c_int64, [c_void_p]),
    "pm_map_query_path": (c_int, [c_void_p]),
    "pm_knn": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "pm_knn_static": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "pm_gather": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_void_p]),
    "pm_gather_radius": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "pm_hit_order_create": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "pm_hit_order_destroy": (None, [c_void_p]),
    "pm_gather_radius_ordered": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "pm_view_photons": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_float, c_int, c_int,
                                c_void_p]),
    "pm_view_matrix": (None, [c_void_p, c_void_p, c_void_p, c_float, c_int, c_int, c_void_p]),
    "pm_render": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
         c_int64, c_int64, c_void_p],
    ),
    "pm_render_rgb": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
         c_int64, c_int64, c_void_p, c_void_p],
    ),
    "pm_selftest_math": (c_int, [c_void_p, c_int64, c_uint32, POINTER(c_int64), POINTER(c_int64), POINTER(c_int64)]),
    "pm_photons_write_text": (c_int, [c_char_p, c_void_p, c_int64]),
    "pm_photons_read_text": (c_int, [c_char_p, c_void_p, c_int64, POINTER(c_int64)]),
    "pm_photons_write_bin": (c_int, [c_char_p, c_void_p, c_int64]),
    "pm_photons_read_bin": (c_int, [c_char_p, c_void_p, c_int64, POINTER(c_int64)]),
}
ABI_SYMBOLS = tuple(_SIGNATURES)

_lib = None


class PmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libpmhot error {code}: {msg}")
        self.code = code


def lib() -> ctypes.CDLL:
    """Load libpmhot.so (built in-tree by ``__graft_entry__.build()``); raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)"
            )
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library does not export the ABI
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(code: int) -> None:
    if code != 0:
        raise PmError(code, lib().pm_last_error().decode("utf-8", "replace"))


def _ptr(x) -> int | None:
    """Raw address of a numpy array (host) or torch tensor (host or device); None stays None."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        if not x.flags["C_CONTIGUOUS"]:
            raise ValueError("numpy arrays handed to libpmhot must be C-contiguous")
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        if not x.is_contiguous():
            raise ValueError("tensors handed to libpmhot must be contiguous")
        return x.data_ptr()
    if isinstance(x, int):
        return x
    raise TypeError(f"cannot take the address of {type(x)!r}")


class Context:
    """pm_ctx: one CUDA device + one stream (include/pmhot.h)."""

    def __init__(self, device: int = 0):
        h = c_void_p()
        _check(lib().pm_ctx_create(device, ctypes.byref(h)))
        self.handle = h
        self.device = device

    def close(self) -> None:
        if getattr(self, "handle", None):
            lib().pm_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self) -> None:
        _check(lib().pm_ctx_sync(self.handle))

    def after_torch(self, *buffers) -> None:
        """Order the context's (non-blocking) stream after torch's current stream when any of `buffers` is a CUDA
        tensor: torch ops that produced it may still be queued (pmhot.h: input readiness). No host wait."""
        for b in buffers:
            if b is not None and getattr(b, "is_cuda", False):
                import torch

                s = torch.cuda.current_stream(b.device)
                if s.cuda_stream != self.stream:
                    _check(lib().pm_ctx_wait_stream(self.handle, s.cuda_stream))
                return

    @property
    def stream(self) -> int:
        return lib().pm_ctx_stream(self.handle) or 0

    @property
    def launch_count(self) -> int:
        return lib().pm_ctx_launch_count(self.handle)

    def reset_launch_count(self) -> None:
        lib().pm_ctx_reset_launch_count(self.handle)

    @property
    def last_kernel_ms(self) -> float:
        return lib().pm_ctx_last_kernel_ms(self.handle)

    @property
    def last_fallback_count(self) -> int:
        return lib().pm_ctx_last_fallback_count(self.handle)

    @property
    def last_handover_count(self) -> int:
        return lib().pm_ctx_last_handover_count(self.handle)


class Comm:
    """pm_comm: this rank's handle on the NCCL communicator the library's multi-GPU calls use (pm_exchange_photons,
    pm_map_build_sharded, pm_ppm_update, pm_comm_allreduce_sum). The reference is single-GPU; SURVEY §8(e)."""

    ID_BYTES = 128

    def __init__(self, ctx: Context, id_bytes: bytes, rank: int, world: int):
        h = c_void_p()
        buf = ctypes.create_string_buffer(bytes(id_bytes), self.ID_BYTES)
        _check(lib().pm_comm_create(ctx.handle, buf, rank, world, ctypes.byref(h)))
        self.ctx, self.handle, self.rank, self.world = ctx, h, rank, world

    @staticmethod
    def unique_id() -> bytes:
        buf = ctypes.create_string_buffer(Comm.ID_BYTES)
        _check(lib().pm_comm_unique_id(buf, Comm.ID_BYTES))
        return buf.raw

    @classmethod
    def from_torch_distributed(cls, ctx: Context, group=None) -> "Comm":
        """Rank 0 draws the NCCL id, torch.distributed (any backend) carries its 128 bytes to the other ranks."""
        import torch
        import torch.distributed as dist

        rank, world = dist.get_rank(group), dist.get_world_size(group)
        dev = torch.device("cuda", ctx.device) if dist.get_backend(group) == "nccl" else torch.device("cpu")
        t = torch.zeros(cls.ID_BYTES, dtype=torch.uint8, device=dev)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(cls.unique_id()), dtype=torch.uint8))
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return cls(ctx, bytes(t.cpu().numpy().tobytes()), rank, world)

    def allreduce_sum(self, t) -> None:
        """In-place sum of a CUDA float32 / int32 / int64 tensor over the ranks."""
        import torch

        code = {torch.float32: 0, torch.int32: 1, torch.int64: 2}[t.dtype]
        self.ctx.after_torch(t)
        _check(lib().pm_comm_allreduce_sum(self.ctx.handle, self.handle, t.data_ptr(), t.numel(), code))

    def close(self) -> None:
        if getattr(self, "handle", None):
            lib().pm_comm_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PhotonMap:
    """pm_map: a photon array in left-balanced kd-tree order (≙ cukd::buildTree output)."""

    def __init__(self, ctx: Context, handle: c_void_p, n: int, bounds=None, keepalive=None):
        self.ctx, self.handle, self.n, self.bounds = ctx, handle, n, bounds
        self._keepalive = keepalive

    @classmethod
    def build(cls, ctx: Context, photons) -> "PhotonMap":
        """In-place build (≙ ray-tracer/src/hostCode.cu:94). photons: PHOTON44 numpy array (host) or a
        torch uint8/any CUDA tensor holding n*44 bytes."""
        n = _count(photons, 44)
        h = c_void_p()
        bounds = np.zeros(6, np.float32)
        ctx.after_torch(photons)
        _check(lib().pm_map_build(ctx.handle, _ptr(photons), n, bounds.ctypes.data, ctypes.byref(h)))
        return cls(ctx, h, n, bounds, photons)

    @classmethod
    def from_tree(cls, ctx: Context, photons) -> "PhotonMap":
        n = _count(photons, 44)
        h = c_void_p()
        ctx.after_torch(photons)
        _check(lib().pm_map_from_tree(ctx.handle, _ptr(photons), n, ctypes.byref(h)))
        return cls(ctx, h, n, None, photons)

    @classmethod
    def build_sharded(cls, ctx: Context, photons, rank: int, world: int, reduce_sum) -> "PhotonMap":
        """pm_map_build_shard + pm_map_build_finish. photons: CUDA torch tensor of n*44 bytes, identical on every rank
        (reordered in place); `reduce_sum(t)` sums an int32 CUDA tensor over the ranks in place (e.g.
        ``lambda t: dist.all_reduce(t)``). The map and the array equal PhotonMap.build's byte for byte."""
        import torch

        n = _count(photons, 44)
        elem = torch.empty(n, dtype=torch.int32, device=photons.device)
        dim = torch.zeros((n + 3) // 4 * 4, dtype=torch.uint8, device=photons.device)
        bounds = np.zeros(6, np.float32)
        ctx.after_torch(photons)  # also orders the zero fill of `dim` before the library's writes
        _check(lib().pm_map_build_shard(ctx.handle, _ptr(photons), n, rank, world, elem.data_ptr(), dim.data_ptr(),
                                        bounds.ctypes.data))
        if world > 1:
            reduce_sum(elem)
            reduce_sum(dim.view(torch.int32))  # one rank contributes each byte: no carries
        h = c_void_p()
        ctx.after_torch(photons)  # the all-reduces run on torch's stream
        _check(lib().pm_map_build_finish(ctx.handle, _ptr(photons), n, elem.data_ptr(), dim.data_ptr(), bounds.ctypes.data,
                                         ctypes.byref(h)))
        return cls(ctx, h, n, bounds, photons)

    @classmethod
    def build_sharded_comm(cls, ctx: Context, comm: "Comm | None", photons) -> "PhotonMap":
        """pm_map_build_sharded: shard pass + the two node-table all-reduces + finish, all inside the library."""
        n = _count(photons, 44)
        h = c_void_p()
        bounds = np.zeros(6, np.float32)
        ctx.after_torch(photons)
        _check(lib().pm_map_build_sharded(ctx.handle, comm.handle if comm else None, _ptr(photons), n, bounds.ctypes.data,
                                          ctypes.byref(h)))
        return cls(ctx, h, n, bounds, photons)

    @classmethod
    def build_cells(cls, ctx: Context, photons, cell_size: float) -> "PhotonMap":
        """pm_map_build_cells: cells-only map (input order, no kd-tree) for `gather_radius`; kNN calls reject it."""
        n = _count(photons, 44)
        h = c_void_p()
        ctx.after_torch(photons)
        _check(lib().pm_map_build_cells(ctx.handle, _ptr(photons), n, cell_size, ctypes.byref(h)))
        return cls(ctx, h, n, None, photons)

    def close(self) -> None:
        if getattr(self, "handle", None):
            lib().pm_map_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def knn(self, queries, k: int, cutoff: float = 100.0, out_ids=None, out_dist2=None, out_r2=None, want_dist2=True,
            want_r2=True):
        """≙ cukd::stackBased::knn<HeapCandidateList<k>>. Returns (ids[nq,k], dist2[nq,k], r2[nq])."""
        nq = _count(queries, 12)
        if isinstance(queries, np.ndarray):
            out_ids = np.empty((nq, k), np.int32) if out_ids is None else out_ids
            if want_dist2 and out_dist2 is None:
                out_dist2 = np.empty((nq, k), np.float32)
            if want_r2 and out_r2 is None:
                out_r2 = np.empty(nq, np.float32)
        else:
            import torch

            dev = queries.device
            out_ids = torch.empty((nq, k), dtype=torch.int32, device=dev) if out_ids is None else out_ids
            if want_dist2 and out_dist2 is None:
                out_dist2 = torch.empty((nq, k), dtype=torch.float32, device=dev)
            if want_r2 and out_r2 is None:
                out_r2 = torch.empty(nq, dtype=torch.float32, device=dev)
        self.ctx.after_torch(queries, out_ids)
        _check(lib().pm_knn(self.ctx.handle, self.handle, _ptr(queries), nq, k, cutoff, _ptr(out_ids), _ptr(out_dist2),
                            _ptr(out_r2)))
        return out_ids, out_dist2, out_r2

    def knn_static(self, queries, k: int, cutoff: float, out_ids, out_dist2=None, out_r2=None) -> None:
        """pm_knn_static: pm_knn on CUDA tensors the caller owns, with the library's resident scratch array — the call
        allocates nothing (no reference counterpart; see pmhot.h). At most KNN_STATIC_MAX_QUERIES queries."""
        self.ctx.after_torch(queries, out_ids)
        _check(lib().pm_knn_static(self.ctx.handle, self.handle, _ptr(queries), _count(queries, 12), k, cutoff, _ptr(out_ids),
                                   _ptr(out_dist2), _ptr(out_r2)))

    def gather(self, hit, brdf, k: int = 50, cutoff: float = 100.0, out_rgb=None):
        """≙ gatherPhotons (ray-tracer/cuda/shading.h:93-121). Returns rgb[nq,3]."""
        nq = _count(hit, 12)
        if out_rgb is None:
            if isinstance(hit, np.ndarray):
                out_rgb = np.empty((nq, 3), np.float32)
            else:
                import torch

                out_rgb = torch.empty((nq, 3), dtype=torch.float32, device=hit.device)
        self.ctx.after_torch(hit, brdf, out_rgb)
        _check(lib().pm_gather(self.ctx.handle, self.handle, _ptr(hit), _ptr(brdf), nq, k, cutoff, _ptr(out_rgb)))
        return out_rgb


class HitOrder:
    """pm_hit_order: the spatial order of a fixed set of hit points, made once and reused by every progressive pass
    (`PhotonMap.gather_radius(..., order=...)` with the same hit / brdf arrays)."""

    def __init__(self, ctx: "Context", hit, brdf):
        self.ctx, self.nq = ctx, _count(hit, 12)
        h = c_void_p()
        ctx.after_torch(hit, brdf)
        _check(lib().pm_hit_order_create(ctx.handle, _ptr(hit), _ptr(brdf), self.nq, ctypes.byref(h)))
        self.handle = h

    def close(self) -> None:
        if getattr(self, "handle", None):
            lib().pm_hit_order_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _gather_radius(self, hit, brdf, radius2, out_flux=None, out_count=None, order: "HitOrder | None" = None):
    """pm_gather_radius / pm_gather_radius_ordered: (flux[nq,3], count[nq]) for per-point squared radii (one
    progressive pass)."""
    nq = _count(hit, 12)
    if isinstance(hit, np.ndarray):
        out_flux = np.empty((nq, 3), np.float32) if out_flux is None else out_flux
        out_count = np.empty(nq, np.int32) if out_count is None else out_count
    else:
        import torch

        out_flux = torch.empty((nq, 3), dtype=torch.float32, device=hit.device) if out_flux is None else out_flux
        out_count = torch.empty(nq, dtype=torch.int32, device=hit.device) if out_count is None else out_count
    self.ctx.after_torch(hit, brdf, radius2, out_flux)
    if order is not None:
        _check(lib().pm_gather_radius_ordered(self.ctx.handle, self.handle, order.handle, _ptr(hit), _ptr(brdf), _ptr(radius2), nq,
                                              _ptr(out_flux), _ptr(out_count)))
    else:
        _check(lib().pm_gather_radius(self.ctx.handle, self.handle, _ptr(hit), _ptr(brdf), _ptr(radius2), nq, _ptr(out_flux),
                                      _ptr(out_count)))
    return out_flux, out_count


PhotonMap.gather_radius = _gather_radius


def _count(x, rec_bytes: int) -> int:
    if isinstance(x, np.ndarray):
        nbytes = x.nbytes
    else:
        nbytes = x.numel() * x.element_size()
    if nbytes % rec_bytes:
        raise ValueError(f"buffer of {nbytes} bytes is not a whole number of {rec_bytes}-byte records")
    return nbytes // rec_bytes


class Scene:
    """pm_scene: triangle soup + per-mesh materials with an LBVH (≙ loadGeometry, common/src/world.cpp:3-58)."""

    def __init__(self, ctx: Context, verts: np.ndarray, tris: np.ndarray, tri_mesh: np.ndarray, mats: np.ndarray):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 3)
        tris = np.ascontiguousarray(tris, np.int32).reshape(-1, 3)
        tri_mesh = np.ascontiguousarray(tri_mesh, np.int32).reshape(-1)
        mats = np.ascontiguousarray(mats, MATERIAL).reshape(-1)
        h = c_void_p()
        _check(lib().pm_scene_create(ctx.handle, verts.ctypes.data, len(verts), tris.ctypes.data, len(tris),
                                     tri_mesh.ctypes.data, mats.ctypes.data, len(mats), ctypes.byref(h)))
        self.ctx, self.handle = ctx, h
        self.nt = len(tris)

    def close(self) -> None:
        if getattr(self, "handle", None):
            lib().pm_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def trace_closest(self, rays: np.ndarray) -> np.ndarray:
        rays = np.ascontiguousarray(rays, RAY)
        hits = np.empty(len(rays), HIT)
        _check(lib().pm_trace_closest(self.ctx.handle, self.handle, rays.ctypes.data, len(rays), hits.ctypes.data))
        return hits

    def view_photons(self, photons: np.ndarray, look_from, look_at, look_up, fovy: float, width: int, height: int) -> np.ndarray:
        """pm_view_photons (≙ photonViewer): RGBA8 image (height, width) of the photons visible from the camera."""
        photons = np.ascontiguousarray(photons, PHOTON40).reshape(-1)
        cam = [np.ascontiguousarray(v, np.float32).reshape(3) for v in (look_from, look_at, look_up)]
        fb = np.empty((height, width), np.uint32)
        _check(lib().pm_view_photons(self.ctx.handle, self.handle, photons.ctypes.data, len(photons), cam[0].ctypes.data,
                                     cam[1].ctypes.data, cam[2].ctypes.data, float(fovy), width, height, fb.ctypes.data))
        return fb

    def trace_occluded(self, rays: np.ndarray) -> np.ndarray:
        rays = np.ascontiguousarray(rays, RAY)
        occ = np.empty(len(rays), np.uint8)
        _check(lib().pm_trace_occluded(self.ctx.handle, self.handle, rays.ctypes.data, len(rays), occ.ctypes.data))
        return occ

    def emit_photons_keyed(self, lights: np.ndarray, casted: int, max_depth: int, caustics: bool, seed: int = 0,
                           rng_mode: int = RNG_LCG_PARITY, first_photon: int = 0, photon_stride: int = 1):
        """pm_emit_photons_keyed on host buffers: (photons[:count], keys[:count] uint64, n_paths, n_segments)."""
        lights = np.ascontiguousarray(lights, LIGHT)
        capacity = (int(casted) // max(1, photon_stride) + len(lights) + 1) * max(1, int(max_depth))
        out = np.empty(capacity, PHOTON40)
        keys = np.empty(capacity, np.uint64)
        count, npaths, nseg = c_int64(0), c_int64(0), c_int64(0)
        _check(lib().pm_emit_photons_keyed(self.ctx.handle, self.handle, lights.ctypes.data, len(lights), casted, max_depth,
                                           1 if caustics else 0, seed, rng_mode, first_photon, photon_stride,
                                           out.ctypes.data, keys.ctypes.data, capacity, ctypes.byref(count),
                                           ctypes.byref(npaths), ctypes.byref(nseg)))
        return out[: count.value], keys[: count.value], npaths.value, nseg.value

    def emit_photons(self, lights: np.ndarray, casted: int, max_depth: int, caustics: bool, seed: int = 0,
                     rng_mode: int = RNG_LCG_PARITY, first_photon: int = 0, photon_stride: int = 1, out=None,
                     capacity: int | None = None, shapes: np.ndarray | None = None):
        """≙ runNormal / runCaustics (photon-mapping/src/hostCode.cu:112-138).
        Returns (photons[:count], n_paths, n_segments). `shapes` (LIGHT_SHAPE per light) selects square emitters."""
        lights = np.ascontiguousarray(lights, LIGHT)
        if capacity is None:
            capacity = int(casted) * int(max_depth)  # hostCode.cu:93
        if out is None:
            out = np.empty(capacity, PHOTON40)
        count, npaths, nseg = c_int64(0), c_int64(0), c_int64(0)
        self.ctx.after_torch(out)
        if shapes is not None:
            shapes = np.ascontiguousarray(shapes, LIGHT_SHAPE).reshape(-1)
            assert len(shapes) == len(lights)
            _check(lib().pm_emit_photons_shaped(self.ctx.handle, self.handle, lights.ctypes.data, shapes.ctypes.data, len(lights),
                                                casted, max_depth, 1 if caustics else 0, seed, rng_mode, first_photon,
                                                photon_stride, _ptr(out), None, capacity, ctypes.byref(count),
                                                ctypes.byref(npaths), ctypes.byref(nseg)))
            res = out[: count.value] if isinstance(out, np.ndarray) else out
            return res, count.value, npaths.value, nseg.value
        _check(lib().pm_emit_photons(self.ctx.handle, self.handle, lights.ctypes.data, len(lights), casted, max_depth,
                                     1 if caustics else 0, seed, rng_mode, first_photon, photon_stride, _ptr(out),
                                     capacity, ctypes.byref(count), ctypes.byref(npaths), ctypes.byref(nseg)))
        res = out[: count.value] if isinstance(out, np.ndarray) else out
        return res, count.value, npaths.value, nseg.value

    def render(self, lights: np.ndarray, global_map: PhotonMap | None, caustic_map: PhotonMap | None, camera: np.ndarray,
               width: int, height: int, spp: int = 1, depth: int = 1, sky=(1.0, 1.0, 1.0), pixel_begin: int = 0,
               pixel_end: int | None = None, linear: bool = False):
        """≙ owlRayGenLaunch2D(simpleRayGen) (ray-tracer/src/hostCode.cu:231-237). Returns uint32[H,W]; with
        ``linear`` also the colours before make_rgba as float32[W*H, 3] by launch index (pm_render_rgb)."""
        lights = np.ascontiguousarray(lights, LIGHT)
        camera = np.ascontiguousarray(camera, CAMERA)
        sky = np.asarray(sky, np.float32)
        fb = np.zeros((height, width), np.uint32)
        if pixel_end is None:
            pixel_end = width * height
        rgb = np.zeros((width * height, 3), np.float32) if linear else None
        _check(lib().pm_render_rgb(self.ctx.handle, self.handle, lights.ctypes.data, len(lights),
                                   global_map.handle if global_map else None, caustic_map.handle if caustic_map else None,
                                   camera.ctypes.data, width, height, spp, depth, sky.ctypes.data, pixel_begin, pixel_end,
                                   fb.ctypes.data, rgb.ctypes.data if linear else None))
        return (fb, rgb) if linear else fb


def photons_to_maps(ctx: Context, global40: np.ndarray, caustic40: np.ndarray):
    """≙ loadPhotons' record conversion (ray-tracer/src/hostCode.cu:65-83). Returns (global44, caustic44)."""
    g = np.ascontiguousarray(global40, PHOTON40)
    c = np.ascontiguousarray(caustic40, PHOTON40)
    og = np.zeros(len(g) + len(c), PHOTON44)
    oc = np.zeros(len(c), PHOTON44)
    _check(lib().pm_photons_to_maps(ctx.handle, g.ctypes.data if len(g) else None, len(g),
                                    c.ctypes.data if len(c) else None, len(c),
                                    og.ctypes.data if len(og) else None, oc.ctypes.data if len(oc) else None))
    return og, oc


def make_camera(look_from, look_at, look_up, fovy: float, width: int, height: int) -> np.ndarray:
    """setupCamera (ray-tracer/src/hostCode.cu:100-108), float32 arithmetic as written there.
    Note the reference scales the image plane by cos(fovy), not tan(fovy/2)."""
    f32 = np.float32
    frm, at, up = (np.asarray(v, f32) for v in (look_from, look_at, look_up))

    def normalize(v):
        d = f32(f32(v[0] * v[0] + v[1] * v[1]) + v[2] * v[2])
        return (v * f32(f32(1.0) / np.sqrt(d, dtype=f32))).astype(f32)

    def cross(a, b):
        return np.array([a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1]], f32)

    aspect = f32(width / f32(height))
    cos_fovy = f32(np.cos(f32(fovy), dtype=f32))
    cam = np.zeros((), CAMERA)
    d00 = normalize(at - frm)
    du = (f32(cos_fovy * aspect) * normalize(cross(d00, up))).astype(f32)
    dv = (cos_fovy * normalize(cross(du, d00))).astype(f32)
    d00 = (d00 - f32(0.5) * (du + dv)).astype(f32)
    cam["pos"], cam["dir_00"], cam["dir_du"], cam["dir_dv"] = frm, d00, du, dv
    return cam


def write_photons_text(path: str, photons: np.ndarray) -> None:
    p = np.ascontiguousarray(photons, PHOTON40)
    _check(lib().pm_photons_write_text(path.encode(), p.ctypes.data, len(p)))


def read_photons_text(path: str) -> np.ndarray:
    cnt = c_int64(0)
    _check(lib().pm_photons_read_text(path.encode(), None, 0, ctypes.byref(cnt)))
    out = np.zeros(cnt.value, PHOTON40)
    if cnt.value:
        _check(lib().pm_photons_read_text(path.encode(), out.ctypes.data, cnt.value, ctypes.byref(cnt)))
    return out


def write_photons_bin(path: str, photons: np.ndarray) -> None:
    p = np.ascontiguousarray(photons, PHOTON40)
    _check(lib().pm_photons_write_bin(path.encode(), p.ctypes.data, len(p)))


def read_photons_bin(path: str) -> np.ndarray:
    cnt = c_int64(0)
    _check(lib().pm_photons_read_bin(path.encode(), None, 0, ctypes.byref(cnt)))
    out = np.zeros(cnt.value, PHOTON40)
    if cnt.value:
        _check(lib().pm_photons_read_bin(path.encode(), out.ctypes.data, cnt.value, ctypes.byref(cnt)))
    return out
