"""Multi-GPU partitioning of the inference path: pure host logic, no collective on the data path.

* frame-pair sharding (BASELINE configs[3]): rank r of g takes pairs ``pair_range(r, g, n)``; weights are
  replicated; nothing is exchanged.
* spatial tiling with halo (BASELINE configs[4], one 4K pair over 8 GPUs): ``tile_plan`` cuts the /16-padded frame
  into a grid of tiles, each extended by a halo and snapped to multiples of 16 (the encoder strides by 2 four
  times, models/DAT.py:20-29 double back up); a rank runs the whole forward on its tile+halo crop with the GLOBAL
  rgb mean (models/BaseModel.py:14-17 reduces over the full frames) and keeps the interior.

The reference has neither: it scales only by DistributedDataParallel over batches (Trainer.py:19).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

__all__ = ["pair_range", "Tile", "tile_plan", "tiles_of_rank", "default_grid"]


def pair_range(rank: int, world: int, n_pairs: int) -> Tuple[int, int]:
    """Contiguous block partition of ``n_pairs`` frame pairs: [begin, end) of ``rank``; sizes differ by at most 1."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(n_pairs, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


@dataclass(frozen=True)
class Tile:
    """Crop [y0:y1, x0:x1] of the padded frame is computed; [iy0:iy1, ix0:ix1] (frame coordinates) is kept."""
    y0: int
    y1: int
    x0: int
    x1: int
    iy0: int
    iy1: int
    ix0: int
    ix1: int

    @property
    def crop(self):
        return slice(self.y0, self.y1), slice(self.x0, self.x1)

    @property
    def interior_in_crop(self):
        return slice(self.iy0 - self.y0, self.iy1 - self.y0), slice(self.ix0 - self.x0, self.ix1 - self.x0)

    @property
    def interior(self):
        return slice(self.iy0, self.iy1), slice(self.ix0, self.ix1)


def _cuts(n: int, parts: int, divisor: int) -> List[int]:
    """parts+1 cut positions over [0, n], interior cuts snapped to multiples of ``divisor``."""
    cuts = [0]
    for i in range(1, parts):
        c = int(round(n * i / parts / divisor)) * divisor
        cuts.append(min(max(c, cuts[-1] + divisor), n))
    cuts.append(n)
    return cuts


def tile_plan(H: int, W: int, grid: Tuple[int, int], halo: int, divisor: int = 16) -> List[Tile]:
    """Row-major list of ``grid[0] x grid[1]`` tiles covering an H x W frame (H, W multiples of ``divisor``)."""
    if H % divisor or W % divisor:
        raise ValueError(f"frame {H}x{W} must be padded to a multiple of {divisor} first")
    gy, gx = grid
    if gy < 1 or gx < 1 or gy * divisor > H or gx * divisor > W:
        raise ValueError(f"grid {grid} does not fit a {H}x{W} frame")
    halo = -(-halo // divisor) * divisor
    ys, xs = _cuts(H, gy, divisor), _cuts(W, gx, divisor)
    tiles = []
    for j in range(gy):
        for i in range(gx):
            iy0, iy1, ix0, ix1 = ys[j], ys[j + 1], xs[i], xs[i + 1]
            tiles.append(Tile(max(0, iy0 - halo), min(H, iy1 + halo), max(0, ix0 - halo), min(W, ix1 + halo),
                              iy0, iy1, ix0, ix1))
    return tiles


def tiles_of_rank(tiles: List[Tile], rank: int, world: int) -> List[Tile]:
    b, e = pair_range(rank, world, len(tiles))
    return tiles[b:e]


def default_grid(world: int) -> Tuple[int, int]:
    """2 x 4 for 8 ranks, otherwise the most square factorisation with more columns than rows."""
    best = (1, world)
    for gy in range(1, world + 1):
        if world % gy == 0 and gy <= world // gy:
            best = (gy, world // gy)
    return best
