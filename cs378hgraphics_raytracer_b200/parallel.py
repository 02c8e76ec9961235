"""Multi-GPU plumbing: one process per GPU, scene replicated, image split into interleaved 32x32
tiles (tile t belongs to rank t % world), NCCL only for the scene broadcast and the framebuffer
gather (BASELINE.json north_star).  The pixel work needs no collective at all: every
pixel is independent (reference RayTracer.cpp:283-313).

torch.distributed is plumbing here; nothing in this file computes pixels.  The same functions run
on the gloo backend with CPU tensors (tests/test_parallel_cpu.py)."""
import numpy as np
import torch
import torch.distributed as dist

TILE = 32


def tile_grid(w, h):
    return (w + TILE - 1) // TILE, (h + TILE - 1) // TILE


def tiles_of_rank(w, h, rank, world):
    tx, ty = tile_grid(w, h)
    return list(range(rank, tx * ty, world))


class _DevPtr:
    """Wraps a raw device pointer (the frame rt_render_device leaves in HBM) for torch.as_tensor."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def device_frame(ptr, w, h, device):
    return torch.as_tensor(_DevPtr(ptr, w * h * 3), device=device).view(h, w, 3)


def pack_tiles(frame, rank, world):
    """frame: (h, w, 3) uint8 tensor -> (n_own_tiles, TILE*TILE*3) of this rank's tiles, in tile order."""
    h, w, _ = frame.shape
    tx, ty = tile_grid(w, h)
    padded = frame.new_zeros((ty * TILE, tx * TILE, 3))
    padded[:h, :w] = frame
    tiles = padded.view(ty, TILE, tx, TILE, 3).permute(0, 2, 1, 3, 4).reshape(tx * ty, TILE * TILE * 3)
    return tiles[rank::world].contiguous()


def unpack_tiles(gathered, w, h, world):
    """gathered: list (per rank) of (n_tiles_of_rank, TILE*TILE*3) -> (h, w, 3) frame."""
    tx, ty = tile_grid(w, h)
    tiles = gathered[0].new_zeros((tx * ty, TILE * TILE * 3))
    for r in range(world):
        n = len(range(r, tx * ty, world))
        tiles[r::world] = gathered[r][:n]
    full = tiles.view(ty, tx, TILE, TILE, 3).permute(0, 2, 1, 3, 4).reshape(ty * TILE, tx * TILE, 3)
    return full[:h, :w].contiguous()


def gather_frame(frame, rank, world, dst=0):
    """Framebuffer gather: every rank contributes its packed tiles (equal, padded counts) with one
    all_gather; rank `dst` un-interleaves.  Returns the full frame on dst, None elsewhere."""
    h, w, _ = frame.shape
    mine = pack_tiles(frame, rank, world)
    if world == 1:
        return unpack_tiles([mine], w, h, 1)
    tx, ty = tile_grid(w, h)
    cap = (tx * ty + world - 1) // world
    send = mine.new_zeros((cap, TILE * TILE * 3))
    send[:mine.shape[0]] = mine
    recv = mine.new_empty((world, cap, TILE * TILE * 3))
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1))
    if rank != dst:
        return None
    return unpack_tiles([recv[r] for r in range(world)], w, h, world)


_gather_bufs = {}


def gather_frame_device(ctx, ptr, w, h, rank, world, device, dst=0):
    """Framebuffer gather of the device frame rt_render_device left at `ptr`: the rank's tiles are packed by a device kernel
    (rt_pack_tiles), ONE gather sends each rank's block to rank `dst` only (NCCL send / recv; 1 / world of the frame per
    rank instead of the whole frame to everyone), rank dst un-interleaves with a device kernel (rt_unpack_tiles).
    torch only carries the bytes.  Returns the (h, w, 3) device frame on dst, None elsewhere; buffers are reused."""
    tx, ty = tile_grid(w, h)
    cap = (tx * ty + world - 1) // world
    key = (w, h, world, str(device))
    if key not in _gather_bufs:
        packed = torch.zeros((cap, TILE * TILE * 3), dtype=torch.uint8, device=device)
        big = torch.empty((world, cap, TILE * TILE * 3), dtype=torch.uint8, device=device) if rank == dst else None
        full = torch.empty((h, w, 3), dtype=torch.uint8, device=device) if rank == dst else None
        _gather_bufs[key] = (packed, big, full)
    packed, big, full = _gather_bufs[key]
    ctx.pack_tiles(int(ptr), w, h, rank, world, packed.data_ptr())
    if world > 1:
        torch.cuda.current_stream(device).synchronize()  # the library's stream finished the pack (the call is synchronous); order torch's stream after it
        dist.gather(packed, list(big.unbind(0)) if rank == dst else None, dst=dst)
    elif rank == dst:
        big[0].copy_(packed)
    if rank != dst:
        return None
    torch.cuda.current_stream(device).synchronize()
    ctx.unpack_tiles(big.data_ptr(), w, h, world, cap, full.data_ptr())
    return full


def broadcast_arena(arena_bytes, rank, world, device, src=0):
    """Scene broadcast: rank `src` holds the flattened arena (host bytes); everyone gets a copy."""
    if world == 1:
        return arena_bytes
    n = torch.tensor([len(arena_bytes) if rank == src else 0], dtype=torch.int64, device=device)
    dist.broadcast(n, src)
    if rank == src:
        buf = torch.frombuffer(bytearray(arena_bytes), dtype=torch.uint8).to(device)
    else:
        buf = torch.empty(int(n.item()), dtype=torch.uint8, device=device)
    dist.broadcast(buf, src)
    return buf.cpu().numpy().tobytes()


def upload_bytes(ctx, arena_bytes):
    """rt_upload_scene from a Python bytes object."""
    import ctypes
    from .binding import lib
    buf = np.frombuffer(arena_bytes, dtype=np.uint8)
    ctx._keep = buf
    ctx._check(lib().rt_upload_scene(ctx._h, ctypes.c_void_p(buf.ctypes.data), len(arena_bytes)))
