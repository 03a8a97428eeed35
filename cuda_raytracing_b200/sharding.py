"""How the path shards across the GPUs of one box; pure host logic, no CUDA here.

Sample sharding (north_star's primary): rank g renders its own samples of EVERY pixel into an f32 SUM buffer
with XORWOW subsequence = pixel + g*n_pixels (rank 0 therefore reproduces the reference stream exactly); the
buffers are summed with one NCCL reduce over NVLink and scaled by 1/total_spp in the resolve kernel.
Tile sharding: rank g renders interleaved bands of rows with the reference stream (bit-identical to 1 GPU).
"""


def sample_shard(rank, world, n_pixels, total_spp=None, per_rank_spp=None):
    """-> (subsequence_offset, spp_for_this_rank, total_spp).  Weak scaling: per_rank_spp fixed; strong: total_spp split
    with the first total_spp % world ranks taking one extra sample."""
    if rank < 0 or rank >= world:
        raise ValueError("rank out of range")
    if per_rank_spp is not None:
        return rank * n_pixels, per_rank_spp, per_rank_spp * world
    base, extra = divmod(total_spp, world)
    return rank * n_pixels, base + (1 if rank < extra else 0), total_spp


def tile_rows(rank, world, height, band=8):
    """Interleaved bands of `band` rows: rank g owns bands g, g+world, ...; returns a list of (row_begin,row_end)."""
    out = []
    b = rank
    while b * band < height:
        out.append((b * band, min(height, (b + 1) * band)))
        b += world
    return out


def reduce_scale(total_spp):
    """Scale applied once, after the reduce, by the resolve kernel (f32 like the reference's pixel_samples_scale)."""
    import numpy as np
    return float(np.float32(1.0) / np.float32(total_spp))
