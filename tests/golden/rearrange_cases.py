"""Case list and seeded inputs shared by make_rearrange_vectors.py (which runs the reference) and the tests. No side
effects on import."""
import numpy as np

# (name, shape [N, B, H, C], d) for the bare kernel, both directions
KERNEL_CASES = [("k_d1", (2, 8, 8, 8), 1), ("k_d2", (2, 8, 8, 8), 2), ("k_d4", (1, 16, 4, 16), 4),
                ("k_d8", (2, 4, 8, 16), 8), ("k_70b_tp4_d2", (2, 64, 2, 128), 2), ("k_70b_tp2_d4", (1, 64, 4, 128), 4)]
# connector-level: prefill rank `rank` of TP `ptp` writes to / reads from a decode engine of TP ptp*d
#   (name, layers, N local, N remote, B, H local, C, d, local ids, staging ids, remote ids)
CONNECTOR_CASES = [
    ("c_d2", 2, 12, 9, 16, 4, 16, 2, [1, 2, 3, 7], [8, 9, 10, 11], [5, 0, 3, 8]),
    ("c_d4", 1, 12, 6, 16, 4, 16, 4, [0, 5, 6], [9, 10, 11], [2, 4, 1]),
    ("c_d1", 2, 8, 8, 16, 4, 16, 1, [6, 1], [3, 4], [0, 7]),
    ("c_d2_split_ranges", 1, 16, 10, 16, 2, 32, 2, [0, 1, 2, 8, 9], [11, 12, 14, 15, 5], [9, 3, 4, 0, 6]),
]


def case_input(seed: int, shape) -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.integers(-2 ** 15, 2 ** 15, size=shape, dtype=np.int64).astype(np.int16)


def connector_inputs(idx: int, layers, n_local, n_remote, B, H, C, d):
    local = [case_input(1000 * idx + l, (2, n_local, B, H, C)) for l in range(layers)]
    remote = [[case_input(1000 * idx + 100 * (g + 1) + l, (2, n_remote, B, H // d, C)) for l in range(layers)]
              for g in range(d)]
    return local, remote
