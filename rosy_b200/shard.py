"""Per-texture sharding of an asset's texture set over ranks.

The reference walks the images of an asset serially on one thread (src/Packager/Gltf.cpp:209-285).
Textures are independent (one file in, one file out, Packager.cpp:216-308), so the set shards over
GPUs with no collective: inside one process the C-ABI work queue (rosy_tex_submit / wait_all) feeds
every configured device; across processes (one rank per GPU, bench.py under torchrun) each rank takes
the slice computed here.  The plan is a pure function of the item sizes, so every rank computes the
same plan without talking to the others, and the output bytes do not depend on the rank count.
"""
from typing import List, Sequence, Tuple


def dedupe(image_indices: Sequence[int]) -> List[int]:
    """First occurrence of every image index, in order.  The reference's dedupe compares
    (image_index, ordinal) tuples that never match (Gltf.cpp:127,205-209: F14), so an image shared by
    N materials is compressed N times; compressing it once produces the same files."""
    seen, out = set(), []
    for i in image_indices:
        if i not in seen:
            seen.add(i)
            out.append(i)
    return out


def plan(costs: Sequence[int], world: int) -> List[List[int]]:
    """Longest-processing-time-first assignment of items (cost = level-0 pixel count) to `world`
    ranks.  Deterministic: ties go to the lower item index, then the lower rank."""
    if world < 1:
        raise ValueError("world must be >= 1")
    order = sorted(range(len(costs)), key=lambda i: (-int(costs[i]), i))
    load = [0] * world
    shards: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        shards[r].append(i)
        load[r] += int(costs[i])
    return shards


def my_items(costs: Sequence[int], rank: int, world: int) -> List[int]:
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    return plan(costs, world)[rank]


def imbalance(costs: Sequence[int], world: int) -> Tuple[int, float]:
    """(max load, max load / mean load) of the plan: the bound on weak-scaling efficiency."""
    shards = plan(costs, world)
    loads = [sum(int(costs[i]) for i in s) for s in shards]
    mean = sum(loads) / float(world) if world else 0.0
    return max(loads) if loads else 0, (max(loads) / mean if mean > 0 else 1.0)
