"""Device-resident containers that stand in for the reference's numpy state
vector and scipy block-diagonal matrices.

* ``DeviceVector``  -- the interleaved state ``x[n_params*i + p]``
  (kafka/inference/utils.py:201) kept planar on the GPU; converts to the
  reference layout on ``np.asarray`` / indexing.
* ``BlockDiag``     -- the block-diagonal (inverse) covariance the reference
  builds with ``block_diag`` (kafka/inference/utils.py:240-339), kept as packed
  lower triangles; supports what callers of the reference use: ``.diagonal()``
  (input_output/observations.py:393, kf_tools.py:302), ``.dot`` (kf_tools.py:90),
  ``+`` (kf_tools.py:89), ``*1.`` (linear_kf.py:238), ``.shape``, ``.tocsc()``.
* ``PixelIndex``    -- state-mask compaction (linear_kf.py:79-80).

All arithmetic runs in libkafka_b200.so; torch only owns the memory.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

ALIGN = 32            # pixels; planes start 256-byte aligned


def padded(n):
    return max(ALIGN, (int(n) + ALIGN - 1) // ALIGN * ALIGN)


def tri_count(p):
    return p * (p + 1) // 2


def default_device():
    if not torch.cuda.is_available():
        raise _lib.KafkaCudaError("no CUDA device: kafka_inferenceengine_b200 has no CPU path")
    return torch.device("cuda", torch.cuda.current_device())


def _empty(rows, ld, device, dtype=torch.float64):
    return torch.empty((rows, ld), dtype=dtype, device=device)


def rows_zero_padded(rows, n_pix, device, dtype=torch.float64):
    """(rows, padded(n_pix)) buffer for kernels that write every pixel < n_pix of every
    row: only the padding columns are cleared (a full-size fill is 0.16 ms per array at
    15 M pixels, three to four arrays per date)."""
    ld = padded(n_pix)
    t = torch.empty((rows, ld), dtype=dtype, device=device)
    if ld > n_pix:
        t[:, n_pix:].zero_()
    return t


class PixelIndex(object):
    """Compaction of a boolean state mask: ``idx[k]`` = row-major raster index of
    the k-th True pixel (bit-exact with ``np.flatnonzero(mask.ravel())``)."""

    def __init__(self, state_mask, device=None):
        lib = _lib.get_lib()
        device = device or default_device()
        self.device = device
        mask = np.ascontiguousarray(np.asarray(state_mask) != 0).astype(np.uint8)
        self.shape = mask.shape
        self.n_raster = int(mask.size)
        self.mask_dev = torch.from_numpy(mask.ravel()).to(device, non_blocking=False)
        scratch = torch.empty(int(lib.kafka_compact_scratch_elems(self.n_raster)), dtype=torch.int32, device=device)
        n_true = C.c_int64(0)
        with torch.cuda.device(device):
            _lib.check(lib.kafka_compact_mask(_lib.ptr(self.mask_dev), self.n_raster, None, 0, _lib.ptr(scratch),
                                              C.byref(n_true), _lib.stream_ptr()), "kafka_compact_mask(count)")
            self.n_pix = int(n_true.value)
            self.idx = torch.empty(max(self.n_pix, 1), dtype=torch.int32, device=device)
            _lib.check(lib.kafka_compact_mask(_lib.ptr(self.mask_dev), self.n_raster, _lib.ptr(self.idx),
                                              self.n_pix, _lib.ptr(scratch), C.byref(n_true),
                                              _lib.stream_ptr()), "kafka_compact_mask")
        self.ld = padded(self.n_pix)

    def gather_band(self, obs, inv_var, obs_mask, y, w, m):
        """Device rasters -> compacted y/w/m rows (solvers.py:90-93)."""
        lib = _lib.get_lib()
        with torch.cuda.device(self.device):
            _lib.check(lib.kafka_gather_band(_lib.ptr(obs), _lib.ptr(inv_var), _lib.ptr(obs_mask),
                                             _lib.ptr(self.idx), self.n_pix, _lib.ptr(y), _lib.ptr(w),
                                             _lib.ptr(m), _lib.stream_ptr()), "kafka_gather_band")

    def gather_bands(self, obs, inv_var, obs_mask, n_bands, y, w, m):
        """All bands of one date in one launch: (n_bands, n_raster) device rasters ->
        (n_bands, ld) compacted rows."""
        lib = _lib.get_lib()
        with torch.cuda.device(self.device):
            _lib.check(lib.kafka_gather_bands(_lib.ptr(obs), _lib.ptr(inv_var), _lib.ptr(obs_mask), obs.stride(0),
                                              n_bands, _lib.ptr(self.idx), self.n_pix, y.stride(0), _lib.ptr(y),
                                              _lib.ptr(w), _lib.ptr(m), _lib.stream_ptr()), "kafka_gather_bands")

    def scatter_output(self, x, a_packed, n_params, want_unc=True):
        """(n_params, ny, nx) float32 rasters of the mean and 1/sqrt(diag(P^-1))
        (input_output/observations.py:374-375,392-393)."""
        lib = _lib.get_lib()
        # one allocation for both rasters (views of it are handed out): one fill instead of two, and the
        # output stage reads both back with one copy
        both = torch.zeros((2 if want_unc else 1, n_params, self.n_raster), dtype=torch.float32, device=self.device)
        mean_r = both[0]
        unc_r = both[1] if want_unc else None
        with torch.cuda.device(self.device):
            _lib.check(lib.kafka_scatter_output(_lib.ptr(x.data), _lib.ptr(a_packed.data) if want_unc else None,
                                                x.ld, self.n_pix, n_params, _lib.ptr(self.idx), self.n_raster,
                                                _lib.ptr(mean_r), _lib.ptr(unc_r), _lib.stream_ptr()),
                       "kafka_scatter_output")
        shape = (n_params,) + tuple(self.shape)
        return mean_r.view(shape), (unc_r.view(shape) if want_unc else None)


class DeviceVector(object):
    """State vector of ``n_pix`` pixels x ``n_params`` parameters, planar on the GPU."""

    __array_priority__ = 1000

    def __init__(self, data, n_pix, n_params):
        self.data = data              # torch (n_params, ld) float64 cuda
        self.n_pix = int(n_pix)
        self.n_params = int(n_params)
        self._host = None
        self._uniform = None          # None = unknown, False = pixels differ, ndarray = the shared vector

    # -- construction -----------------------------------------------------
    @classmethod
    def empty(cls, n_pix, n_params, device=None):
        device = device or default_device()
        return cls(_empty(n_params, padded(n_pix), device), n_pix, n_params)

    def uniform_value(self):
        """The state vector shared by ALL pixels, or None when pixels differ.
        This is the test behind the reference's "one unique vector" shortcut
        (kafka/inference/utils.py:72-74); decided once per vector by a device
        reduction (kafka_state_uniform) and cached -- device vectors are never
        modified after they are handed out."""
        if self._uniform is None:
            if self.n_pix == 0:
                self._uniform = False
            else:
                # [pixel 0's vector | n_diff (i32) + padding], read back in one kafka_read_small
                tmp = torch.zeros(self.n_params + 1, dtype=torch.float64, device=self.device)
                n_diff = tmp[self.n_params:].view(torch.int32)[0:1]
                with torch.cuda.device(self.device):
                    _lib.check(_lib.get_lib().kafka_state_uniform(_lib.ptr(self.data), self.ld, self.n_pix,
                                                                  self.n_params, _lib.ptr(n_diff), _lib.stream_ptr()),
                               "kafka_state_uniform")
                    tmp[:self.n_params].copy_(self.data[:, 0])
                    host = _lib.read_small(tmp)
                self._uniform = host[:self.n_params].copy() if int(host[self.n_params:].view(np.int32)[0]) == 0 \
                    else False
        return None if self._uniform is False else self._uniform

    @classmethod
    def from_interleaved(cls, x, n_params, device=None):
        device = device or default_device()
        if isinstance(x, DeviceVector):
            return x
        if isinstance(x, torch.Tensor):
            t = x.to(device=device, dtype=torch.float64).contiguous().view(-1)
        else:
            a = np.ascontiguousarray(x, dtype=np.float64).ravel()
            if not a.flags.writeable:                     # e.g. a slice of an .npz: torch wants a writable source
                a = a.copy()
            t = torch.from_numpy(a).to(device)
        if t.numel() % n_params:
            raise ValueError("state vector length %d is not a multiple of n_params=%d" % (t.numel(), n_params))
        n_pix = t.numel() // n_params
        out = cls.empty(n_pix, n_params, device)
        with torch.cuda.device(device):
            _lib.check(_lib.get_lib().kafka_interleaved_to_planar(_lib.ptr(t), n_pix, n_params, _lib.ptr(out.data),
                                                                  out.ld, _lib.stream_ptr()),
                       "kafka_interleaved_to_planar")
        return out

    @classmethod
    def broadcast(cls, mean, n_pix, device=None):
        device = device or default_device()
        host = np.array(mean.detach().cpu().numpy() if isinstance(mean, torch.Tensor) else mean,
                        dtype=np.float64).ravel()
        mean = torch.as_tensor(host, device=device)
        out = cls.empty(n_pix, mean.numel(), device)
        out.data.copy_(mean[:, None].expand(-1, out.ld))
        out._uniform = host                      # known without asking the device
        return out

    # -- properties ---------------------------------------------------------
    @property
    def ld(self):
        return self.data.shape[1]

    @property
    def device(self):
        return self.data.device

    @property
    def shape(self):
        return (self.n_pix * self.n_params,)

    @property
    def dtype(self):
        return np.dtype(np.float64)

    def __len__(self):
        return self.n_pix * self.n_params

    # -- conversion -----------------------------------------------------------
    def interleaved(self):
        """Device tensor in the reference layout x[n_params*i+p]."""
        out = torch.empty(self.n_pix * self.n_params, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().kafka_planar_to_interleaved(_lib.ptr(self.data), self.ld, self.n_pix,
                                                                  self.n_params, _lib.ptr(out), _lib.stream_ptr()),
                       "kafka_planar_to_interleaved")
        return out

    def numpy(self):
        if self._host is None:
            self._host = self.interleaved().cpu().numpy()
        return self._host

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a if dtype is None else a.astype(dtype, copy=False)

    def __getitem__(self, key):
        return self.numpy()[key]

    def astype(self, dtype):
        return self.numpy().astype(dtype)

    def copy(self):
        return DeviceVector(self.data.clone(), self.n_pix, self.n_params)

    def __mul__(self, s):
        if isinstance(s, (int, float)):
            return DeviceVector(self.data * float(s), self.n_pix, self.n_params)
        return self.numpy() * s

    __rmul__ = __mul__

    def __sub__(self, other):
        return self.numpy() - np.asarray(other)

    def __add__(self, other):
        return self.numpy() + np.asarray(other)


class BlockDiag(object):
    """Block-diagonal symmetric matrix, one ``n_params`` x ``n_params`` block per
    pixel, stored as packed lower triangles (planar).  ``broadcast`` blocks hold a
    single block shared by every pixel (e.g. a prior that is the same everywhere,
    kafka_test_S2.py:106-118)."""

    def __init__(self, data, n_pix, n_params, broadcast=False):
        self.data = data              # (L, ld) or (L,) when broadcast
        self.n_pix = int(n_pix)
        self.n_params = int(n_params)
        self.is_broadcast = bool(broadcast)

    # -- construction -----------------------------------------------------
    @classmethod
    def empty(cls, n_pix, n_params, device=None):
        device = device or default_device()
        return cls(_empty(tri_count(n_params), padded(n_pix), device), n_pix, n_params)

    @classmethod
    def from_block(cls, block, n_pix, device=None):
        """One block repeated for every pixel (broadcast storage)."""
        device = device or default_device()
        block = np.asarray(block, dtype=np.float64)
        P = block.shape[0]
        r, c = np.tril_indices(P)
        return cls(torch.from_numpy(np.ascontiguousarray(block[r, c])).to(device), n_pix, P, broadcast=True)

    @classmethod
    def from_blocks(cls, blocks, device=None):
        """Dense (n_pix, P, P) blocks (numpy or torch)."""
        device = device or default_device()
        if isinstance(blocks, torch.Tensor):
            t = blocks.to(device=device, dtype=torch.float64).contiguous()
        else:
            t = torch.from_numpy(np.ascontiguousarray(blocks, dtype=np.float64)).to(device)
        n_pix, P, _ = t.shape
        out = cls.empty(n_pix, P, device)
        with torch.cuda.device(device):
            _lib.check(_lib.get_lib().kafka_blocks_to_packed(_lib.ptr(t), n_pix, P, _lib.ptr(out.data), out.ld,
                                                             _lib.stream_ptr()), "kafka_blocks_to_packed")
        return out

    @classmethod
    def from_scipy(cls, A, n_params, device=None):
        """A scipy sparse block-diagonal matrix as built by the reference's
        ``block_diag`` (utils.py:240-339)."""
        import scipy.sparse as sp
        A = sp.bsr_matrix(A, blocksize=(n_params, n_params))
        n_pix = A.shape[0] // n_params
        A.sort_indices()
        rows = np.repeat(np.arange(n_pix), np.diff(A.indptr))
        if not np.array_equal(rows, A.indices):
            raise ValueError("matrix is not block diagonal with %dx%d blocks" % (n_params, n_params))
        blocks = np.zeros((n_pix, n_params, n_params))
        blocks[rows] = A.data
        return cls.from_blocks(blocks, device)

    @classmethod
    def coerce(cls, A, n_params, n_pix, device=None):
        if A is None or isinstance(A, BlockDiag):
            return A
        if isinstance(A, np.ndarray) and A.ndim == 3:
            return cls.from_blocks(A, device)
        if isinstance(A, np.ndarray) and A.ndim == 2 and A.shape == (n_params, n_params):
            return cls.from_block(A, n_pix, device)
        return cls.from_scipy(A, n_params, device)

    # -- properties ---------------------------------------------------------
    @property
    def device(self):
        return self.data.device

    @property
    def ld(self):
        return padded(self.n_pix) if self.is_broadcast else self.data.shape[1]

    @property
    def shape(self):
        n = self.n_pix * self.n_params
        return (n, n)

    def materialise(self):
        """Per-pixel storage (expands a broadcast block)."""
        if not self.is_broadcast:
            return self
        data = self.data[:, None].expand(-1, padded(self.n_pix)).contiguous()
        return BlockDiag(data, self.n_pix, self.n_params)

    # -- what the reference's callers use -------------------------------------
    def diagonal(self):
        """Main diagonal, interleaved like the state vector (numpy)."""
        return self.diagonal_device().cpu().numpy()

    def diagonal_device(self):
        m = self.materialise()
        out = torch.empty(self.n_pix * self.n_params, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().kafka_packed_diagonal(_lib.ptr(m.data), m.ld, self.n_pix, self.n_params,
                                                            _lib.ptr(out), _lib.stream_ptr()),
                       "kafka_packed_diagonal")
        return out

    def blocks_device(self):
        m = self.materialise()
        out = torch.empty((self.n_pix, self.n_params, self.n_params), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().kafka_packed_to_blocks(_lib.ptr(m.data), m.ld, self.n_pix, self.n_params,
                                                             _lib.ptr(out), _lib.stream_ptr()),
                       "kafka_packed_to_blocks")
        return out

    def blocks(self):
        return self.blocks_device().cpu().numpy()

    def to_scipy(self, dtype=None):
        import scipy.sparse as sp
        b = self.blocks()
        if dtype is not None:
            b = b.astype(dtype)
        n = self.n_pix
        return sp.bsr_matrix((b, np.arange(n), np.arange(n + 1)), shape=self.shape)

    def tocsc(self):
        return self.to_scipy().tocsc()

    def tocsr(self):
        return self.to_scipy().tocsr()

    def todense(self):
        return self.to_scipy().todense()

    def toarray(self):
        return self.to_scipy().toarray()

    def dot(self, vec):
        """Block-diagonal matrix-vector product (kf_tools.py:90) -> numpy."""
        v = DeviceVector.from_interleaved(vec, self.n_params, self.device)
        x = v.interleaved().view(self.n_pix, self.n_params, 1)
        y = torch.bmm(self.blocks_device(), x).view(-1)      # boundary convenience, not on the hot path
        return y.cpu().numpy()

    def copy(self):
        return BlockDiag(self.data.clone(), self.n_pix, self.n_params, self.is_broadcast)

    def __mul__(self, s):
        if not isinstance(s, (int, float)):
            return NotImplemented
        return BlockDiag(self.data * float(s), self.n_pix, self.n_params, self.is_broadcast)

    __rmul__ = __mul__

    def _binary(self, other, sign):
        other = BlockDiag.coerce(other, self.n_params, self.n_pix, self.device)
        if other.n_pix != self.n_pix or other.n_params != self.n_params:
            raise ValueError("block-diagonal shapes differ")
        if self.is_broadcast and other.is_broadcast:
            return BlockDiag(self.data + sign * other.data, self.n_pix, self.n_params, True)
        a, b = self.materialise(), other.materialise()
        return BlockDiag(a.data + sign * b.data, self.n_pix, self.n_params)

    def __add__(self, other):
        if isinstance(other, (int, float)) and other == 0:
            return self
        return self._binary(other, 1.0)

    __radd__ = __add__

    def __sub__(self, other):
        if isinstance(other, (int, float)) and other == 0:
            return self
        return self._binary(other, -1.0)


class DeviceBands(object):
    """Per-band vectors (innovations, forward model) kept band-major like the
    reference's ``np.hstack`` over bands (solvers.py:119-122)."""

    def __init__(self, data, n_pix):
        self.data = data              # (n_bands, ld)
        self.n_pix = int(n_pix)

    def __array__(self, dtype=None, copy=None):
        a = self.data[:, :self.n_pix].contiguous().view(-1).cpu().numpy()
        return a if dtype is None else a.astype(dtype, copy=False)

    def __len__(self):
        return self.data.shape[0] * self.n_pix
