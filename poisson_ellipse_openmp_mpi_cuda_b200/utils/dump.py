"""Solution dump.  The reference never writes its solution; the format is ours
(section 7.3): the full ``(M+1) x (N+1)`` nodal field including the zero Dirichlet boundary, row = i (x),
column = j (y).

* ``*.bin``: one ASCII header line ``poisson_b200 M N float64 little-endian row=i(x) col=j(y)\\n`` followed
  by ``(M+1)*(N+1)`` raw little-endian doubles.
* ``*.npy``: numpy array.
* anything else: text ``x y u`` triples, a blank line between i-rows (gnuplot ``splot`` / ``numpy.loadtxt``).
"""
from __future__ import annotations

import numpy as np
import torch

HEADER = "poisson_b200 {M} {N} float64 little-endian row=i(x) col=j(y)\n"


def node_coordinates(M: int, N: int):
    x = -1.0 + np.arange(M + 1) * (2.0 / M)
    y = -0.6 + np.arange(N + 1) * (1.2 / N)
    return x, y


def write_solution(path: str, w: torch.Tensor, M: int, N: int) -> None:
    arr = w.detach().cpu().numpy().astype("<f8", copy=False)
    assert arr.shape == (M + 1, N + 1)
    if path.endswith(".bin"):
        with open(path, "wb") as f:
            f.write(HEADER.format(M=M, N=N).encode())
            f.write(np.ascontiguousarray(arr).tobytes())
    elif path.endswith(".npy"):
        np.save(path, arr)
    else:
        x, y = node_coordinates(M, N)
        with open(path, "w") as f:
            for i in range(M + 1):
                block = np.column_stack([np.full(N + 1, x[i]), y, arr[i]])
                np.savetxt(f, block, fmt="%.17g")
                f.write("\n")


def read_solution(path: str):
    """Returns ``(M, N, array)`` for any of the three formats."""
    if path.endswith(".bin"):
        with open(path, "rb") as f:
            head = f.readline().decode().split()
            M, N = int(head[1]), int(head[2])
            arr = np.frombuffer(f.read(), dtype="<f8").reshape(M + 1, N + 1)
        return M, N, arr
    if path.endswith(".npy"):
        arr = np.load(path)
        return arr.shape[0] - 1, arr.shape[1] - 1, arr
    data = np.loadtxt(path)
    ys = data[:, 1]
    n1 = int(np.argmax(ys[1:] <= ys[0])) + 1 if len(ys) > 1 and np.any(ys[1:] <= ys[0]) else len(ys)
    arr = data[:, 2].reshape(-1, n1)
    return arr.shape[0] - 1, arr.shape[1] - 1, arr
