"""Maps: cell-name grids -> uint8 cell codes -> the device map pool.

The reference builds every map on the host as a numpy array of cell names
(``GameMap.grid``, mapgen/types.py:8-9) and hands ``grid.tolist()`` to the core
(simulator/simulator.py:70-73, 264-267). Here the same name grids are lowered once to cell codes
(index into the CELLTYPES section of the lowered config), pre-split into per-map object tables,
and kept in HBM so that auto-reset is a device-side copy.

``arena_maps`` is a synthetic generator with the *shape* of the reference arena map builder
(MapGen 2x2 rooms of 25x25, outer border 6, each room filled by scenes/random.py:19-70 with
wall 10, assembler 5, mine_red 10, generator_red 5, lasery 1, armory 1 and 6 agents;
builder/envs.py:57-75). It is not bit-identical to numpy-seeded MapGen output and does not need
to be: maps are inputs to both the simulator and the oracle.
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .layout import L
from .lowering import LoweredConfig

EMPTY_NAMES = ("empty", ".", " ")  # mettagrid_c.cpp:187


def names_to_codes(lowered: LoweredConfig, grid) -> np.ndarray:
    """list[list[str]] / ndarray of names -> uint8 codes; unknown names raise like mettagrid_c.cpp:191-193."""
    arr = np.asarray(grid)
    out = np.zeros(arr.shape, dtype=np.uint8)
    for name in np.unique(arr):
        s = str(name)
        if s in EMPTY_NAMES:
            continue
        if s not in lowered.cell_codes:
            raise RuntimeError(f"Unknown object type: {s}")
        out[arr == name] = lowered.cell_codes[s]
    return out


def codes_to_names(lowered: LoweredConfig, codes: np.ndarray) -> list[list[str]]:
    table = {0: "empty"}
    table.update(lowered.cell_names)
    return [[table[int(v)] for v in row] for row in codes]


@dataclass
class MapPool:
    """Device-ready pool (host copies). Grid cell encoding, u16:
    0 empty | code (wall types) | 0x4000+agent | 0x8000+assembler instance | 0xC000+chest instance."""

    codes: np.ndarray  # [M,H,W] u8
    grids: np.ndarray  # [M,H*W] u16
    agents: np.ndarray  # [M,A,3] i32 (r, c, cell code)
    n_asm: np.ndarray  # [M] i32
    asms: np.ndarray  # [M,max_asm,3] i32 (r, c, cell code)
    n_chest: np.ndarray  # [M] i32
    chests: np.ndarray  # [M,max_chest,3] i32

    @property
    def max_assemblers(self) -> int:
        return int(self.asms.shape[1])

    @property
    def max_chests(self) -> int:
        return int(self.chests.shape[1])


GRID_AGENT, GRID_ASM, GRID_CHEST = 0x4000, 0x8000, 0xC000


def build_pool(lowered: LoweredConfig, codes: np.ndarray) -> MapPool:
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    if codes.ndim == 2:
        codes = codes[None]
    M, H, W = codes.shape
    A = lowered.num_agents
    kind_of = np.zeros(256, dtype=np.uint8)
    for code, kind in lowered.cell_kind.items():
        kind_of[code] = kind
    if int(codes.max(initial=0)) >= int(lowered.words[L.MGC_NUM_CELLTYPES]):
        raise RuntimeError("map uses a cell code the config does not define")
    kinds = kind_of[codes.reshape(M, -1)]
    is_agent, is_asm, is_chest = kinds == L.MG_KIND_AGENT, kinds == L.MG_KIND_ASSEMBLER, kinds == L.MG_KIND_CHEST
    n_agents = is_agent.sum(1)
    if not np.all(n_agents == A):
        bad = int(np.argmax(n_agents != A))
        raise RuntimeError(f"map {bad} has {int(n_agents[bad])} agents, config says {A}")
    n_asm, n_chest = is_asm.sum(1).astype(np.int32), is_chest.sum(1).astype(np.int32)
    max_asm, max_chest = max(1, int(n_asm.max())), max(1, int(n_chest.max()))
    if max_asm >= 0x4000 or max_chest >= 0x4000 or A >= 0x4000:
        raise RuntimeError("too many objects of one kind for the u16 grid encoding")
    flat = codes.reshape(M, -1)
    grids = flat.astype(np.uint16)
    agents = np.zeros((M, A, 3), dtype=np.int32)
    asms = np.zeros((M, max_asm, 3), dtype=np.int32)
    chests = np.zeros((M, max_chest, 3), dtype=np.int32)

    def fill(mask, base, table):
        # row-major rank of each object among its kind = creation order (mettagrid_c.cpp:181-244)
        rank = np.cumsum(mask, axis=1) - 1
        m_idx, cell = np.nonzero(mask)
        k = rank[m_idx, cell]
        grids[m_idx, cell] = (base + k).astype(np.uint16)
        table[m_idx, k, 0] = cell // W
        table[m_idx, k, 1] = cell % W
        table[m_idx, k, 2] = flat[m_idx, cell]

    fill(is_agent, GRID_AGENT, agents)
    fill(is_asm, GRID_ASM, asms)
    fill(is_chest, GRID_CHEST, chests)
    return MapPool(codes=codes, grids=grids, agents=agents, n_asm=n_asm, asms=asms, n_chest=n_chest, chests=chests)


def arena_maps(lowered: LoweredConfig, num_maps: int, seed: int = 0, num_agents: int = 24) -> np.ndarray:
    """Synthetic arena-shaped maps as cell codes [M,62,62]. See module docstring."""
    rooms = num_agents // 6
    rows = int(np.ceil(np.sqrt(rooms)))
    cols = int(np.ceil(rooms / rows))
    room, border = 25, 6
    H, W = rows * room + 2 * border, cols * room + 2 * border
    cc = lowered.cell_codes
    wall = cc["wall"]
    fill = [cc["wall"]] * 10 + [cc["assembler"]] * 5 + [cc["mine_red"]] * 10 + [cc["generator_red"]] * 5
    fill += [cc["lasery"]] * 1 + [cc["armory"]] * 1 + [cc["agent.agent"]] * 6
    fill = np.asarray(fill, dtype=np.uint8)
    rng = np.random.default_rng(seed)
    out = np.full((num_maps, H, W), wall, dtype=np.uint8)
    for m in range(num_maps):
        for k in range(rooms):
            r0, c0 = border + (k // cols) * room, border + (k % cols) * room
            cells = np.zeros(room * room, dtype=np.uint8)
            pos = rng.permutation(room * room)[: len(fill)]
            cells[pos] = rng.permutation(fill)
            out[m, r0 : r0 + room, c0 : c0 + room] = cells.reshape(room, room)
        for k in range(rooms, rows * cols):  # unused room slots stay empty floor
            r0, c0 = border + (k // cols) * room, border + (k % cols) * room
            out[m, r0 : r0 + room, c0 : c0 + room] = 0
    return out


def random_maps(lowered: LoweredConfig, num_maps: int, height: int, width: int, counts: dict[str, int], seed: int = 0,
                border: int = 1) -> np.ndarray:
    """Synthetic maps as cell codes [M,H,W]: a wall border and `counts[name]` cells of each named
    object (agents included, e.g. "agent.agent") scattered uniformly inside. Stands in for the
    reference's random scene (mapgen/scenes/random.py:19-70) when a test or benchmark needs many maps."""
    cc = lowered.cell_codes
    fill = np.concatenate([np.full(n, cc[name], dtype=np.uint8) for name, n in counts.items()])
    inner_h, inner_w = height - 2 * border, width - 2 * border
    if len(fill) > inner_h * inner_w:
        raise ValueError("random_maps: more objects than cells")
    rng = np.random.default_rng(seed)
    out = np.full((num_maps, height, width), cc["wall"], dtype=np.uint8)
    for m in range(num_maps):
        cells = np.zeros(inner_h * inner_w, dtype=np.uint8)
        cells[rng.permutation(inner_h * inner_w)[: len(fill)]] = rng.permutation(fill)
        out[m, border : height - border, border : width - border] = cells.reshape(inner_h, inner_w)
    return out


def arena_mapgen_pool(lowered: LoweredConfig, num_maps: int | None = None) -> np.ndarray:
    """The reference's own arena maps (MapGen, builder/envs.py:57-75, map_builder.seed = 1000 + i) as cell
    codes [M,62,62]. The builders are out of scope; 256 of their grids were generated in the build container
    (tests/golden/make_golden.py arena_pool) and are shipped as data."""
    import os

    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "arena_mapgen_pool.npz")
    data = np.load(path)
    names = [str(x) for x in data["names"]]
    table = np.array([0 if n in EMPTY_NAMES else lowered.cell_codes[n] for n in names], dtype=np.uint8)
    codes = table[data["cells"]]
    return codes if num_maps is None else codes[:num_maps]
