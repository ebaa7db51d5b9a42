"""TEST INFRASTRUCTURE (oracle/): counter-based Philox4x32-10 and the injected-uniform contract.

The reference draws from the unseeded global `random.uniform` (main.py:16, utils.py:9), partly
inside forked pool workers, so its output is not reproducible.  Parity therefore
feeds both sides the SAME stream, defined here and restated in pathtracerpython_b200/csrc:

    u(seed, pixel, sample, bounce, dim) = (x >> 8) * 2**-24
    x = word (dim % 4) of Philox4x32-10(key=(seed, 0), ctr=(pixel, sample, bounce, dim // 4))
    uniform(a, b) := a + (b - a) * u                      (random.uniform's own formula)

Dimension map per (pixel, sample, bounce):
    light sample s in {0,1,2}: dim 4s   -> pick_random_triangle      (utils.py:30)
                               dim 4s+1..4s+3 -> sample_bary_coords   (utils.py:23)
    dim 12 -> lobe pick (main.py:240), dim 13 -> u1 for phi (main.py:242),
    dim 14 -> u2 for theta (main.py:243), dim 15 -> spare.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this package.
"""

M0 = 0xD2511F53
M1 = 0xCD9E8D57
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = 0xFFFFFFFF


def philox4x32_10(ctr, key):
    """Philox4x32 with 10 rounds (Salmon et al., SC'11). ctr: 4 u32, key: 2 u32 -> 4 u32."""
    c0, c1, c2, c3 = [int(c) & MASK for c in ctr]
    k0, k1 = [int(k) & MASK for k in key]
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & MASK, p1 & MASK, ((p0 >> 32) ^ c3 ^ k1) & MASK, p0 & MASK
        k0 = (k0 + W0) & MASK
        k1 = (k1 + W1) & MASK
    return c0, c1, c2, c3


def u01(seed, pixel, sample, bounce, dim):
    """The injected uniform in [0, 1): 24 high bits, exact in FP32 and FP64."""
    x = philox4x32_10((pixel, sample, bounce, dim // 4), (seed, 0))[dim % 4]
    return (x >> 8) * (2.0 ** -24)


class Stream:
    """Stateful view of the stream: the driver sets (pixel, sample, bounce) and a starting dim
    before each reference call; every `uniform` call consumes one dimension."""

    def __init__(self, seed):
        self.seed = seed
        self.pixel = self.sample = self.bounce = self.dim = 0
        self.draws = 0

    def at(self, pixel, sample, bounce, dim=0):
        self.pixel, self.sample, self.bounce, self.dim = pixel, sample, bounce, dim

    def uniform(self, a, b):
        u = u01(self.seed, self.pixel, self.sample, self.bounce, self.dim)
        self.dim += 1
        self.draws += 1
        return a + (b - a) * u
