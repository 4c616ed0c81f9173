"""Sharding of a snapshot across the GPUs of one box.

Segments are independent (own min/max, depth, headers, dither streams), so a
snapshot is cut into contiguous particle-ID slabs = contiguous runs of segments,
one per rank, and every rank encodes its slab with no data-path collective.
Only output placement needs an exchange:

  1. all-gather of each rank's compressed byte count (one int64 per rank) ->
     exclusive prefix = the rank's base offset in the global stream;
  2. optionally, gather of the packed stream to one rank (variable sizes:
     point-to-point sends into the sink's buffer at the prefix offsets).

The functions take torch tensors and a process group and never look at the
backend: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
import torch
import torch.distributed as dist


def partition(n_units, world):
    """Contiguous slabs: rank r of `world` gets units [start, start+count)."""
    return [(r * n_units // world, (r + 1) * n_units // world - r * n_units // world) for r in range(world)]


def exchange_sizes(local_bytes, group=None):
    """local_bytes: 1-element int64 tensor (on the device the backend works on).
    Returns (sizes[world], base) as tensors on the same device: every rank's
    compressed size and this rank's exclusive prefix. No host synchronisation."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = torch.empty(world, dtype=torch.int64, device=local_bytes.device)
    dist.all_gather_into_tensor(sizes, local_bytes.reshape(1), group=group)
    base = torch.cumsum(sizes, 0)[rank] - sizes[rank]
    return sizes, base


def global_offsets(local_offsets, base):
    """Field offsets of this rank's slab inside the global stream."""
    return local_offsets + base


def gather_stream(local, sizes, dst=0, group=None):
    """Gather the ranks' packed streams (uint8 tensors of sizes[r] bytes, 16-byte
    multiples) into one contiguous tensor on rank `dst`, in rank order. Returns
    the stream on `dst`, None elsewhere. `sizes` is a host list/1-D tensor."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = [int(s) for s in sizes]
    if rank != dst:
        if sizes[rank]:
            dist.send(local[:sizes[rank]].contiguous(), dst=dst, group=group)
        return None
    out = torch.empty(sum(sizes), dtype=torch.uint8, device=local.device)
    pos = 0
    for r in range(world):
        if r == dst:
            out[pos:pos + sizes[r]].copy_(local[:sizes[r]])
        elif sizes[r]:
            dist.recv(out[pos:pos + sizes[r]], src=r, group=group)
        pos += sizes[r]
    return out


def scatter_stream(stream, sizes, src=0, group=None, device=None):
    """Inverse of gather_stream: rank `src` holds the global stream; every rank
    receives its slab's bytes."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = [int(s) for s in sizes]
    if rank == src:
        pos = 0
        mine = None
        for r in range(world):
            part = stream[pos:pos + sizes[r]]
            if r == src:
                mine = part.clone()
            elif sizes[r]:
                dist.send(part.contiguous(), dst=r, group=group)
            pos += sizes[r]
        return mine
    out = torch.empty(sizes[rank], dtype=torch.uint8, device=device)
    if sizes[rank]:
        dist.recv(out, src=src, group=group)
    return out
