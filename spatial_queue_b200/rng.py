"""Counter-based RNG contract shared by the CUDA tick, the C oracle and the reference shim.

The reference draws `shuffle` (queue_class.py:231, 236) and `random.choices` (:360) from one global Mersenne
Twister in node-visit order, which no parallel implementation can reproduce.  The contract that replaces it
 keys every draw on (seed, tick, node id, draw slot):

    u(seed, t, node, k) = (mix64(seed, t, node, k) >> 11) * 2**-53          in [0, 1)

    slot k = 0   first shuffle of the candidate in-links      -> r = floor(u * n); element r moves to the front,
                 the rest keep their incoming_links order
    slot k = 1   second shuffle at signal nodes               -> same rule on the phase-filtered list
    slot k = 2+j j-th coin toss of this node call             -> move iff not (u * (w0 + cc) < w0), w0 = 1.0 - cc
                 (the exact predicate of CPython's random.choices([0,1], [1-cc, cc]))
"""
M64 = (1 << 64) - 1
C_SEED = 0x243F6A8885A308D3
C1 = 0x9E3779B97F4A7C15
C2 = 0xBF58476D1CE4E5B9
C3 = 0x94D049BB133111EB


def _fmix(x):
    x &= M64
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & M64
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & M64
    x ^= x >> 31
    return x


def mix64(seed, t, node, k):
    h = (seed ^ C_SEED) & M64
    h = _fmix(h + ((t & M64) * C1))
    h = _fmix(h + ((node & M64) * C2))
    h = _fmix(h + ((k & M64) * C3))
    return h


def uniform(seed, t, node, k):
    return (mix64(seed, t, node, k) >> 11) * (1.0 / 9007199254740992.0)
