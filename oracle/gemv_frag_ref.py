"""CPU restatement of the fragment-ordered weight copy the tensor-core GEMV reads (TEST INFRASTRUCTURE - see oracle/__init__.py).

No counterpart in the reference (vLLM keeps ``nn.Linear`` layout, llama.py:164-180): the layout is this build's own
(csrc/linear_gemv.cu ``gemv_frag_kernel`` / ``t3b_repack_weights_gemv``), so the oracle pins it against the only thing that
matters - the dot products the kernel forms from it equal ``x @ W.T``:

    out[tile][ks][half][lane] (8 bf16 each) = W[row(tile, lane >> 2, half)][32 ks + 8 (lane & 3) : + 8]
    row = 16 tile + (lane >> 2) + 8 half                         (plain projection; rows >= N repeat row N - 1)
    row = 8 tile + (lane >> 2)  |  N / 2 + 8 tile + (lane >> 2)  (gate_up: half 0 = gate, half 1 = the matching up feature)

``mma.sync.m16n8k16`` (row-major A, column-major B) lets lane (g, t) = (lane >> 2, lane & 3) supply A[g][2t, 2t+1], A[g+8][2t, 2t+1],
A[g][2t+8, 2t+9], A[g+8][2t+8, 2t+9]; the kernel feeds it elements 0..3 of the lane's 8 (first MMA of a k step) and 4..7 (second),
and reads the activations with the same permutation, so every product pairs W[n][k] with x[r][k] (``mma_emulate``).
"""
from __future__ import annotations

import numpy as np


def frag_order(w: np.ndarray, gate_up: bool = False) -> np.ndarray:
    """w [N, K] (any dtype) -> [tiles, K // 32, 2, 32, 8] in the order of gemv_frag_kernel."""
    N, K = w.shape
    assert K % 32 == 0
    tiles = N // 16 if gate_up else (N + 15) // 16
    out = np.empty((tiles, K // 32, 2, 32, 8), dtype=w.dtype)
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        for half in range(2):
            for tile in range(tiles):
                row = (N // 2 + tile * 8 + g if half else tile * 8 + g) if gate_up else tile * 16 + g + 8 * half
                row = min(row, N - 1)
                out[tile, :, half, lane, :] = w[row].reshape(K // 32, 4, 8)[:, t, :]
    return out


def mma_emulate(frag: np.ndarray, x: np.ndarray, n_out_rows: int, gate_up: bool = False) -> np.ndarray:
    """The products the kernel forms: for every tile / k step / lane the 8 weight elements meet x[r][32 ks + 8 t : + 8].
    frag as float64 [tiles, ksteps, 2, 32, 8], x float64 [R, K] -> y [R, rows covered] (gate rows then up rows for gate_up)."""
    tiles, ksteps = frag.shape[0], frag.shape[1]
    R = x.shape[0]
    xs = x.reshape(R, ksteps, 4, 8)
    y = np.zeros((R, n_out_rows))
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        for half in range(2):
            part = np.einsum("aks,rks->ra", frag[:, :, half, lane, :], xs[:, :, t, :])       # [R, tiles]: this lane's k share
            for tile in range(tiles):
                row = (n_out_rows // 2 + tile * 8 + g if half else tile * 8 + g) if gate_up else tile * 16 + g + 8 * half
                if row < n_out_rows:
                    y[:, row] += part[:, tile]
    return y
