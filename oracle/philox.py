"""Counter-based noise spec shared by the oracle and csrc/philox.cuh (TEST INFRASTRUCTURE, see oracle/__init__.py).

The reference draws eps with tf.random.normal (models.py:13-15), i.e. TensorFlow's own Philox stream, which
cannot be reproduced without TensorFlow.  The replacement therefore defines ITS OWN stream:

  Philox4x32-10 (Salmon et al., SC'11; Random123), key = (seed_lo, seed_hi),
  counter = (row_lo, row_hi, latent_quad, step)     row = global jet index inside the batch,
                                                    latent_quad = l // 4, step = 0 in noise_mode 0
  4 x u32 -> u_k = ((x_k >> 9) + 0.5) * 2^-23  in (0,1)  (exact in fp32)
  (n0,n1) = sqrt(-2 ln u0) * (cos, sin)(2 pi u1),  (n2,n3) likewise from (u2,u3);   eps[row, 4q+j] = n_j

The u32 stream is bit-exact between this file and the CUDA kernel; the normals agree to a few ulp (libm vs
CUDA logf/sincosf), which is why parity tests feed eps explicitly (noise_mode 2).
Pinned against the Random123 known-answer vectors in tests/test_oracle.py.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr, key):
    """ctr: [...,4] uint32, key: [...,2] uint32 (broadcastable) -> [...,4] uint32."""
    c = [np.asarray(ctr[..., i], np.uint32).copy() for i in range(4)]
    k0 = np.asarray(key[..., 0], np.uint32).copy()
    k1 = np.asarray(key[..., 1], np.uint32).copy()
    with np.errstate(over='ignore'):
        for _ in range(10):
            p0 = M0 * c[0].astype(np.uint64)
            p1 = M1 * c[2].astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), (p0 & MASK).astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), (p1 & MASK).astype(np.uint32)
            c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
            k0 = k0 + W0
            k1 = k1 + W1
    return np.stack(c, axis=-1)


def normal_eps(seed, B, L, row_offset=0, step=0, dtype=np.float32):
    """eps[B, L] of the spec above (fp32 arithmetic, as on the device)."""
    nq = (L + 3) // 4
    rows = (np.arange(B, dtype=np.uint64) + np.uint64(row_offset))[:, None]
    q = np.arange(nq, dtype=np.uint32)[None, :]
    ctr = np.empty((B, nq, 4), np.uint32)
    ctr[..., 0] = (rows & MASK).astype(np.uint32)
    ctr[..., 1] = (rows >> np.uint64(32)).astype(np.uint32)
    ctr[..., 2] = q
    ctr[..., 3] = np.uint32(step)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], np.uint32)
    x = philox4x32_10(ctr, key)
    u = ((x >> np.uint32(9)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -23)
    r0 = np.sqrt(np.float32(-2) * np.log(u[..., 0])); t0 = np.float32(2 * np.pi) * u[..., 1]
    r1 = np.sqrt(np.float32(-2) * np.log(u[..., 2])); t1 = np.float32(2 * np.pi) * u[..., 3]
    n = np.stack([r0 * np.cos(t0), r0 * np.sin(t0), r1 * np.cos(t1), r1 * np.sin(t1)], axis=-1)
    return n.reshape(B, nq * 4)[:, :L].astype(dtype)
