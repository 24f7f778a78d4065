"""numpy statement of the integer-only synthetic relay PCM (csrc/synth.h).

Written independently of the C++ header (vectorised uint64 arithmetic) so that tests can
check the host and device generators against it byte for byte.
"""
import numpy as np

SEED = 0x10CA0B2000000001
GAINS_Q15 = np.array([32768, 29491, 26542, 23888, 21499, 19349, 17414, 15673], dtype=np.int64)
_M = (1 << 64) - 1


def _sm64(x):
    x = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def _sm64_int(x):
    x = (x + 0x9E3779B97F4A7C15) & _M
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M
    return x ^ (x >> 31)


def _tone(phase):
    x = (phase & 0x7FFF).astype(np.int64)
    y = (x * (32768 - x)) >> 13
    a = (y * 4000) >> 15
    return np.where(phase & 0x8000, -a, a)


def relay_view(group, relay, n_samples, utt_samples=None, first=0, seed=SEED):
    """int16 samples [first, first+n_samples) of relay `relay` in collision group `group`."""
    utt = n_samples + first if utt_samples is None else utt_samples
    with np.errstate(over="ignore"):
        n = np.arange(first, first + n_samples, dtype=np.uint64)
        gk = _sm64_int(seed ^ ((0x10CA0B20 + group) & _M))
        inc0 = 400 + gk % 1200
        inc1 = 1600 + (gk >> 16) % 1600
        inc2 = 3200 + (gk >> 32) % 2400
        n32 = n.astype(np.int64)
        t = _tone((inc0 * n32) & 0xFFFF) + _tone((inc1 * n32) & 0xFFFF) + _tone((inc2 * n32) & 0xFFFF)
        edge = np.minimum(n32, utt - 1 - n32)
        edge = np.where(n32 >= utt, 0, edge)
        env = np.minimum(edge, 4000) * 8
        rg = _sm64(np.uint64(gk) + n)
        u = ((t * env) >> 15) + ((rg % np.uint64(401)).astype(np.int64) - 200)
        rk = _sm64_int(gk ^ ((0x9E3779B97F4A7C15 * (relay + 1)) & _M))
        rr = _sm64(np.uint64(rk) + n)
        amp = 64 * (relay + 1)
        noise = (rr % np.uint64(2 * amp + 1)).astype(np.int64) - amp
        x = ((u * int(GAINS_Q15[relay & 7])) >> 15) + noise
    return np.clip(x, -32768, 32767).astype(np.int16)


def batch_time_major(group_base, n_streams, group_size, n_msgs, msg_samples, seed=SEED):
    """Same layout as ga_synth_fill_batch_*: out[m, s, k]."""
    utt = n_msgs * msg_samples
    out = np.empty((n_msgs, n_streams, msg_samples), dtype=np.int16)
    for s in range(n_streams):
        x = relay_view(group_base + s // group_size, s % group_size, utt, utt, 0, seed)
        out[:, s, :] = x.reshape(n_msgs, msg_samples)
    return out
