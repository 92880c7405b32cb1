"""2-D owning container with arithmetic (ref include/matrix/matrix.hpp:22-221; the reference uses it in its GEMM
benchmark only).  A thin value-semantics wrapper around a 2-D ``torch.Tensor`` on any device: ``+ - * /`` allocate,
``+= -= *= /=`` work in place, ``A * B`` between matrices is the matrix product (the tcgen05 GEMM on CUDA, the native
AVX GEMM or ATen on the CPU, via ``ops.tensor_ops.gemm``)."""
from __future__ import annotations

import torch


class Matrix:
    def __init__(self, rows: int = 0, cols: int = 0, data=None, device="cpu", dtype=torch.float32):
        if data is not None:
            t = torch.as_tensor(data, dtype=dtype, device=device)
            if t.numel() != rows * cols:
                raise ValueError(f"Matrix: {t.numel()} values for a {rows}x{cols} matrix")
            self._t = t.reshape(rows, cols).clone()
        else:
            self._t = torch.zeros(rows, cols, dtype=dtype, device=device)

    @staticmethod
    def from_tensor(t: torch.Tensor) -> "Matrix":
        if t.dim() != 2:
            raise ValueError("Matrix.from_tensor needs a 2-D tensor")
        m = Matrix.__new__(Matrix)
        m._t = t
        return m

    # -- shape / access ------------------------------------------------------------------------------------------
    def rows(self) -> int:
        return int(self._t.shape[0])

    def cols(self) -> int:
        return int(self._t.shape[1])

    def size(self) -> int:
        return self._t.numel()

    def data(self) -> torch.Tensor:
        return self._t

    def device(self):
        return self._t.device

    def __call__(self, r: int, c: int) -> float:
        return float(self._t[r, c])

    def __getitem__(self, rc):
        return self._t[rc]

    def __setitem__(self, rc, value):
        self._t[rc] = value

    # -- mutation ------------------------------------------------------------------------------------------------
    def fill(self, value: float):
        self._t.fill_(value)

    def fill_random_uniform(self, range_: float, seed=None):
        g = None if seed is None else torch.Generator(device=self._t.device).manual_seed(int(seed))
        self._t.uniform_(-range_, range_, generator=g)

    def fill_random_normal(self, mean: float, stddev: float, seed=None):
        g = None if seed is None else torch.Generator(device=self._t.device).manual_seed(int(seed))
        self._t.normal_(mean, stddev, generator=g)

    def resize(self, rows: int, cols: int):
        """Like the reference: a different shape reallocates (contents are zero), the same shape is a no-op."""
        if (rows, cols) != tuple(self._t.shape):
            self._t = torch.zeros(rows, cols, dtype=self._t.dtype, device=self._t.device)

    def reshape(self, rows: int, cols: int) -> "Matrix":
        if rows * cols != self.size():
            raise ValueError("Total number of elements must remain the same for reshape")
        return Matrix.from_tensor(self._t.reshape(rows, cols).clone())

    def clone(self) -> "Matrix":
        return Matrix.from_tensor(self._t.clone())

    def to(self, device) -> "Matrix":
        return Matrix.from_tensor(self._t.to(device))

    # -- arithmetic ----------------------------------------------------------------------------------------------
    def _same(self, other: "Matrix", what: str):
        if tuple(self._t.shape) != tuple(other._t.shape):
            raise ValueError(f"Matrix dimensions must match for {what}")

    def __add__(self, other: "Matrix") -> "Matrix":
        self._same(other, "addition")
        return Matrix.from_tensor(self._t + other._t)

    def __iadd__(self, other: "Matrix") -> "Matrix":
        self._same(other, "addition")
        self._t += other._t
        return self

    def __sub__(self, other: "Matrix") -> "Matrix":
        self._same(other, "subtraction")
        return Matrix.from_tensor(self._t - other._t)

    def __isub__(self, other: "Matrix") -> "Matrix":
        self._same(other, "subtraction")
        self._t -= other._t
        return self

    def __mul__(self, other):
        if isinstance(other, Matrix):
            if self.cols() != other.rows():
                raise ValueError("Matrix dimensions must match for multiplication")
            from ..ops import tensor_ops
            return Matrix.from_tensor(tensor_ops.gemm(self._t.contiguous(), other._t.contiguous()))
        return Matrix.from_tensor(self._t * float(other))

    __rmul__ = __mul__

    def __imul__(self, scalar: float) -> "Matrix":
        self._t *= float(scalar)
        return self

    def __truediv__(self, scalar: float) -> "Matrix":
        if float(scalar) == 0.0:
            raise ZeroDivisionError("Division by zero")
        return Matrix.from_tensor(self._t / float(scalar))

    def __itruediv__(self, scalar: float) -> "Matrix":
        if float(scalar) == 0.0:
            raise ZeroDivisionError("Division by zero")
        self._t /= float(scalar)
        return self

    def __repr__(self):
        return f"Matrix({self.rows()}x{self.cols()}, device={self._t.device})"
