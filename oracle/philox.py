"""TEST INFRASTRUCTURE ONLY -- numpy restatement of Philox4x32-10 (Salmon, Moraes, Dror, Shaw,
"Parallel random numbers: as easy as 1, 2, 3", SC'11) and of the counter layout the reset kernel
uses (csrc/oige_common.cuh: reset_uniform).  Pinned by the Random123 known-answer vectors in
tests/test_philox.py."""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint32 arrays; returns four uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint32).copy() for c in (c0, c1, c2, c3))
    k0 = np.asarray(k0, dtype=np.uint32).copy()
    k1 = np.asarray(k1, dtype=np.uint32).copy()
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = M0 * c0.astype(np.uint64)
            p1 = M1 * c2.astype(np.uint64)
            n0 = (p1 >> np.uint64(32)).astype(np.uint32) ^ c1 ^ k0
            n1 = (p1 & np.uint64(0xFFFFFFFF)).astype(np.uint32)
            n2 = (p0 >> np.uint64(32)).astype(np.uint32) ^ c3 ^ k1
            n3 = (p0 & np.uint64(0xFFFFFFFF)).astype(np.uint32)
            c0, c1, c2, c3 = n0, n1, n2, n3
            k0 = (k0 + W0).astype(np.uint32)
            k1 = (k1 + W1).astype(np.uint32)
    return c0, c1, c2, c3


def u01(x):
    return ((x >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)


def reset_uniforms(seed: int, counter: int, env_gids, num_dof: int, which: int) -> np.ndarray:
    """[len(env_gids), num_dof] float32 uniforms the kernel draws for (seed, counter, env, dof, which)."""
    gid = np.asarray(env_gids, dtype=np.uint64)
    out = np.zeros((len(gid), num_dof), np.float32)
    c0 = (gid & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    c1 = (gid >> np.uint64(32)).astype(np.uint32)
    c2 = np.full(len(gid), counter & 0xFFFFFFFF, np.uint32)
    k0 = np.full(len(gid), seed & 0xFFFFFFFF, np.uint32)
    k1 = np.full(len(gid), (seed >> 32) & 0xFFFFFFFF, np.uint32)
    hi = ((counter >> 32) & 0xFFFFFFFF) << 8 & 0xFFFFFFFF
    for blk in range((num_dof + 3) // 4):
        c3 = np.full(len(gid), hi ^ (blk * 2 + which), np.uint32)
        r = philox4x32_10(c0, c1, c2, c3, k0, k1)
        for j in range(4):
            d = blk * 4 + j
            if d < num_dof:
                out[:, d] = u01(r[j])
    return out
