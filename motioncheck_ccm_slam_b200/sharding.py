"""Partition arithmetic for the two sharded paths: agent streams onto GPUs, and the server keyframe
database onto ranks. No compute here — the per-shard search and the merge are CUDA kernels (csrc/match_allpairs.cu)."""


def agent_device(agent_id, n_gpus):
    """Agent a -> GPU a mod G (each client's frames on its own GPU; one process / handle / stream per agent)."""
    if n_gpus < 1:
        raise ValueError("n_gpus must be >= 1")
    return int(agent_id) % int(n_gpus)


def shard_range(n_items, rank, world):
    """Contiguous [lo, hi) slice of n_items for `rank` of `world`: sizes differ by at most one, slices tile [0, n)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return n_items * rank // world, n_items * (rank + 1) // world


def keyframe_shards(kf_desc_counts, world):
    """Shard a keyframe database BY KEYFRAME into `world` contiguous keyframe ranges of about equal descriptor counts.
    Returns [(kf_lo, kf_hi, desc_lo, desc_hi)] per rank; desc_lo is the shard's global index base (ccm_best2.idx)."""
    total = sum(kf_desc_counts)
    out, kf, acc = [], 0, 0
    for r in range(world):
        target = total * (r + 1) // world
        kf_lo, d_lo = kf, acc
        while kf < len(kf_desc_counts) and (acc + kf_desc_counts[kf] <= target or r == world - 1):
            acc += kf_desc_counts[kf]
            kf += 1
        out.append((kf_lo, kf, d_lo, acc))
    return out
