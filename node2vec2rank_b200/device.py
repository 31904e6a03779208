"""Device-side operators: thin Python over the C ABI (torch tensors own the memory).

Every function here runs hand-written sm_100a kernels through libn2v2r_b200.so; nothing
falls back to the CPU.
"""
from __future__ import annotations

import ctypes as C
import os
import time
from dataclasses import dataclass, field

import numpy as np
import torch

from . import _native as N


def _dev():
    return N.require_gpu()


@dataclass
class DeviceLayer:
    """A preprocessed graph layer resident in HBM (see include/n2v2r_b200.h): dense bf16 planes, one e4m3 plane, or -- for
    sparse layers -- row lists multiplied entry by entry in fp32 (storage == "rows") or compressed tiles that the product
    kernel expands into dense bf16 tiles for the tensor cores (storage == "tiles")."""
    planes: torch.Tensor | None     # (P, n, ld) bf16, None for a tiles layer
    n: int
    ld: int
    n_planes: int
    symmetric: bool = True
    planes_t: torch.Tensor | None = None
    colsum: torch.Tensor | None = None     # (n,) float64, column sums of the fp32 values
    shape: tuple = field(default=())
    storage: str = "planes"
    tile_rows: int = 0
    tile_ptr: torch.Tensor | None = None   # (n_tiles + 1,) int32 (uint32 offsets)
    entries: torch.Tensor | None = None    # (n_entries, 2) int32: 8-byte entries
    tile_ptr_t: torch.Tensor | None = None
    entries_t: torch.Tensor | None = None
    n_entries: int = 0                     # entries in use (the arrays hold at least one allocation unit)
    n_entries_t: int = 0
    row_offset: int = 0                    # row slab of a layer (make_slab_layer): rows [row_offset, row_offset + rows)
    rows: int = 0                          # 0 = the whole layer
    ld_t: int = 0
    panel_cols: int = 0                    # row lists: columns per panel (tile_rows = rows per block)
    list_cap: int = 0                      # row lists: longest list in bytes (A, A^T)
    list_cap_t: int = 0
    n_units: int = 0                       # row lists: size of the lists in 16-byte units (A, A^T)
    n_units_t: int = 0
    dtype = np.float32

    def __post_init__(self):
        self.shape = (self.n, self.n)

    @property
    def device(self):
        return (self.planes if self.planes is not None else self.tile_ptr).device

    def c_struct(self):
        if self.storage == "rows":
            return N.Layer(0, 0, self.n_planes, N.STORE_ROWS, self.ld,
                           self.tile_ptr.data_ptr(), self.entries.data_ptr(),
                           0 if self.tile_ptr_t is None else self.tile_ptr_t.data_ptr(),
                           0 if self.entries_t is None else self.entries_t.data_ptr(),
                           self.n_entries, self.n_entries_t, self.tile_rows, self.panel_cols, 0, 0, 0, self.list_cap, self.list_cap_t)
        if self.storage == "tiles":
            return N.Layer(0, 0, self.n_planes, N.STORE_TILES, self.ld,
                           self.tile_ptr.data_ptr(), self.entries.data_ptr(),
                           0 if self.tile_ptr_t is None else self.tile_ptr_t.data_ptr(),
                           0 if self.entries_t is None else self.entries_t.data_ptr(),
                           self.n_entries, self.n_entries_t, self.tile_rows, 0, 0, 0, 0)
        return N.Layer(self.planes.data_ptr(), 0 if self.planes_t is None else self.planes_t.data_ptr(),
                       self.n_planes, N.STORE_E4M3 if self.storage == "e4m3" else N.STORE_PLANES, self.ld, 0, 0, 0, 0, 0, 0, 0, 0,
                       self.row_offset, self.rows, self.ld_t)

    def to_dense(self):
        """fp32 (n, n) device tensor reconstructed from the stored form (exact)."""
        if self.storage == "rows":
            out = torch.zeros((self.n, self.n), dtype=torch.float32, device=self.device)
            N.check(N.load().n2v2r_rows_to_dense(N.stream_ptr(), N.ptr(self.tile_ptr), N.ptr(self.entries), self.n, self.tile_rows,
                                                 self.panel_cols, N.ptr(out), self.n))
            return out
        if self.storage == "tiles":
            out = torch.zeros((self.n, self.n), dtype=torch.float32, device=self.device)
            N.check(N.load().n2v2r_tiles_to_dense(N.stream_ptr(), N.ptr(self.tile_ptr), N.ptr(self.entries), self.n, self.tile_rows,
                                                  N.ptr(out), self.n))
            return out
        if self.storage == "e4m3":
            return self.planes[0, :, :self.n].view(torch.float8_e4m3fn).to(torch.float32)
        return self.planes[:, :, :self.n].to(torch.float32).sum(dim=0)      # a row slab gives its (rows, n) block

    def to_e4m3(self):
        """The same layer as ONE e4m3 plane (1 byte per entry, N2V2R_STORE_E4M3): for one-plane layers whose values e4m3
        represents exactly (binarised layers).  Raises ValueError otherwise."""
        if self.storage == "e4m3":
            return self
        if self.storage != "planes" or self.n_planes != 1:
            raise ValueError("only one-plane bf16 layers can be stored as e4m3")
        lib = N.load()
        ld8 = (self.n + 127) // 128 * 128

        def conv(p):
            out = torch.empty((1, self.n, ld8), dtype=torch.uint8, device=p.device)
            exact = C.c_int(0)
            N.check(lib.n2v2r_plane_to_e4m3(N.stream_ptr(), N.ptr(p), self.n, self.n, self.ld, N.ptr(out), ld8, C.byref(exact)))
            if not exact.value:
                raise ValueError("layer values are not exactly representable in e4m3")
            return out

        return DeviceLayer(planes=conv(self.planes), n=self.n, ld=ld8, n_planes=1, symmetric=self.symmetric,
                           planes_t=None if self.planes_t is None else conv(self.planes_t), colsum=self.colsum, storage="e4m3")

    def numpy(self):
        return self.to_dense().cpu().numpy()

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a if dtype is None else a.astype(dtype)

    @property
    def nbytes(self):
        if self.storage == "rows":
            return (self.n_units + self.n_units_t) * 16 + sum(t.numel() * 4 for t in (self.tile_ptr, self.tile_ptr_t) if t is not None)
        if self.storage == "tiles":
            return (self.n_entries + self.n_entries_t) * 8 + sum(t.numel() * 4 for t in (self.tile_ptr, self.tile_ptr_t) if t is not None)
        return self.planes.numel() * self.planes.element_size() * (2 if self.planes_t is not None else 1)


def _dtype_code(t: torch.Tensor):
    if t.dtype == torch.float32:
        return N.F32
    if t.dtype == torch.float64:
        return N.F64
    raise TypeError(f"unsupported dtype {t.dtype}; expected float32 or float64")


def to_device(x):
    """Host array (float32/float64, anything else is cast to float64 like pandas would) -> contiguous device tensor."""
    dev = _dev()
    if isinstance(x, torch.Tensor):
        t = x
    else:
        a = np.asarray(getattr(x, "values", x))
        if a.dtype not in (np.float32, np.float64):
            a = a.astype(np.float64)
        t = torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype not in (torch.float32, torch.float64):
        t = t.to(torch.float64)
    return t.to(dev, non_blocking=True).contiguous()


def to_host(t: torch.Tensor) -> np.ndarray:
    """Device tensor -> NumPy array through pinned memory (torch's caching host allocator): a D2H copy into pageable memory
    is staged at a few GB/s, into pinned memory it runs at PCIe speed.  The array keeps the pinned storage alive."""
    t = t.contiguous()
    try:
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    except RuntimeError:                      # no lockable memory left: the plain pageable copy
        return t.cpu().numpy()
    h.copy_(t, non_blocking=True)
    torch.cuda.current_stream(t.device).synchronize()
    return h.numpy()


def sbm_generate(n, K, B, seed, layer, weighted=False):
    """K0: synthetic SBM layer on the device, fp32 (n, n)."""
    dev = _dev()
    ld = (n + 3) // 4 * 4
    out = torch.empty((n, ld), dtype=torch.float32, device=dev)
    Bh = np.ascontiguousarray(np.asarray(B, dtype=np.float32).reshape(K, K))
    N.check(N.load().n2v2r_sbm_generate(N.stream_ptr(), N.ptr(out), n, ld, K,
                                        Bh.ctypes.data_as(C.POINTER(C.c_float)), C.c_uint64(seed), layer, int(weighted)))
    return out if ld == n else out[:, :n].contiguous()


def dense_from_edges(iu, iv, w, n, dtype=torch.float64):
    """K0b: (n, n) device matrix from a duplicate-free undirected weighted edge list (endpoints as final node positions,
    negative = dropped).  Replaces nx.to_numpy_array(nx.read_weighted_edgelist(...)) (dataloader.py:100-104)."""
    dev = _dev()
    out = torch.zeros((n, n), dtype=dtype, device=dev)
    u = torch.as_tensor(np.ascontiguousarray(iu, dtype=np.int32)).to(dev)
    v = torch.as_tensor(np.ascontiguousarray(iv, dtype=np.int32)).to(dev)
    ww = torch.as_tensor(np.ascontiguousarray(w, dtype=np.float64)).to(dev)
    N.check(N.load().n2v2r_scatter_edges(N.stream_ptr(), N.ptr(u), N.ptr(v), N.ptr(ww), u.numel(), N.ptr(out), _dtype_code(out), n, n))
    return out


CSV_CHUNK_ENTRIES = 16_000_000        # rows are parsed and copied in chunks of about this many cells (128 MB of float64)


def csv_to_device(path, sep, chunk_entries=None):
    """Adjacency file (first column = row labels, first line = column labels; what pd.read_csv(index_col=0, header=0)
    reads, dataloader.py:86-91) -> (float64 (rows, cols) device matrix in FILE order, row labels, column labels).

    The file is parsed by pandas' C reader in chunks of rows; every chunk is copied into one of two pinned staging
    buffers and sent to the device on a copy stream while the next chunk is being parsed: no full float64 DataFrame on
    the host, no pageable copy.  Labels are exactly what pandas infers for the whole file."""
    import pandas as pd
    dev = _dev()
    cols = pd.read_csv(path, index_col=0, header=0, sep=sep, nrows=0).columns
    c = len(cols)
    with open(path, "rb") as f:                       # upper bound of the number of data rows: newlines after the header
        lines = 0
        while True:
            block = f.read(1 << 24)
            if not block:
                break
            lines += block.count(b"\n")
            last = block[-1:]
        if lines and last != b"\n":
            lines += 1
    cap = max(lines - 1, 0)
    out = torch.empty((max(cap, 1), max(c, 1)), dtype=torch.float64, device=dev)
    chunk_rows = max(1, (chunk_entries or CSV_CHUNK_ENTRIES) // max(c, 1))
    stage = [torch.empty((chunk_rows, max(c, 1)), dtype=torch.float64, pin_memory=True) for _ in range(2)]
    sent = [None, None]
    copy = torch.cuda.Stream(dev)
    copy.wait_stream(torch.cuda.current_stream(dev))
    labels, r0, k = [], 0, 0
    for chunk in pd.read_csv(path, index_col=0, header=0, sep=sep, chunksize=chunk_rows):
        vals = chunk.to_numpy(dtype=np.float64)        # ValueError on non-numeric cells
        r = vals.shape[0]
        if r0 + r > cap:
            raise ValueError(f"{path}: more data rows than lines")
        slot = k % 2
        if sent[slot] is not None:
            sent[slot].synchronize()                    # the previous copy out of this staging buffer has finished
        stage[slot][:r].copy_(torch.from_numpy(vals))
        with torch.cuda.stream(copy):
            out[r0:r0 + r].copy_(stage[slot][:r], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy)
        sent[slot] = ev
        labels.append(chunk.index)
        r0 += r
        k += 1
    torch.cuda.current_stream(dev).wait_stream(copy)
    copy.synchronize()                                  # the staging buffers go out of scope
    if not labels:
        rows = pd.Index([])
    elif len({str(ix.dtype) for ix in labels}) == 1 and labels[0].dtype != object:
        rows = labels[0].append(labels[1:]) if len(labels) > 1 else labels[0]
    elif len(labels) == 1:
        rows = labels[0]
    else:            # pandas infers the label type per chunk; mixed chunks -> what it infers for the whole first column
        rows = pd.Index(pd.read_csv(path, usecols=[0], header=0, sep=sep).iloc[:, 0])
    return out[:r0, :c], rows, cols


def gather_submatrix(x: torch.Tensor, row_idx, col_idx, transpose=False):
    """K0d: the (len(row_idx), len(col_idx)) float64 device matrix out[i][j] = x[row_idx[i]][col_idx[j]]
    (transpose: x[col_idx[j]][row_idx[i]]) -- match_networks' reindex (and dataloader.py:103's .T) on the device."""
    lib = N.load()
    if x.dtype != torch.float64 or x.dim() != 2:
        raise TypeError("gather_submatrix expects a float64 matrix")
    ri = torch.as_tensor(np.ascontiguousarray(row_idx, dtype=np.int32)).to(x.device)
    ci = torch.as_tensor(np.ascontiguousarray(col_idx, dtype=np.int32)).to(x.device)
    lim_r, lim_c = (x.shape[1], x.shape[0]) if transpose else (x.shape[0], x.shape[1])
    if ri.numel() and (int(ri.min()) < 0 or int(ri.max()) >= lim_r) or ci.numel() and (int(ci.min()) < 0 or int(ci.max()) >= lim_c):
        raise IndexError("gather_submatrix: index outside the matrix")
    out = torch.empty((ri.numel(), ci.numel()), dtype=torch.float64, device=x.device)
    N.check(lib.n2v2r_gather_submatrix(N.stream_ptr(), N.ptr(x), x.shape[0], x.shape[1], x.stride(0), N.ptr(ri), N.ptr(ci),
                                       ri.numel(), ci.numel(), int(bool(transpose)), N.ptr(out), max(ci.numel(), 1)))
    return out


def select_nonzero(x: torch.Tensor, frac: float, absolute=False, threshold=None):
    """K1a: (nnz, v_lo, v_hi, virtual_index) of the nonzeros of the (abs/thresholded) matrix."""
    lib = N.load()
    ws = torch.empty(lib.n2v2r_select_workspace_bytes(), dtype=torch.uint8, device=x.device)
    nnz, lo, hi, pos = C.c_int64(), C.c_double(), C.c_double(), C.c_double()
    rows, cols = x.shape
    N.check(lib.n2v2r_select_nonzero(N.stream_ptr(), N.ptr(x), _dtype_code(x), rows, cols, cols, int(bool(absolute)),
                                     int(threshold is not None), float(threshold or 0.0), float(frac),
                                     int(x.dtype == torch.float32), N.ptr(ws), ws.numel(),
                                     C.byref(nnz), C.byref(lo), C.byref(hi), C.byref(pos)))
    return nnz.value, lo.value, hi.value, pos.value


def percentile_cut(x: torch.Tensor, q, absolute=False, threshold=None):
    """np.percentile(y[y.nonzero()], q) with NumPy's 'linear' method, the order statistics coming
    from the device radix select (preprocessing_utils.py:134).  The interpolation follows
    numpy/lib/_function_base_impl.py (_quantile/_lerp) in the dtype of the data."""
    ft = np.float32 if x.dtype == torch.float32 else np.float64
    quant = np.true_divide(q, ft(100))                 # percentile(): q / a.dtype.type(100)
    nnz, lo, hi, pos = select_nonzero(x, float(quant), absolute, threshold)
    if nnz == 0:
        raise IndexError("index -1 is out of bounds for axis 0 with size 0")   # what np.percentile raises on empty input
    virt = ft(pos)
    prev = np.floor(virt)
    gamma = ft(np.float64(virt) - prev)
    a, b, t = ft(lo), ft(hi), gamma
    diff = b - a
    lerp = a + diff * t
    if t >= 0.5:
        lerp = b - diff * (1 - t)
    return lerp


def planes_needed(x32: torch.Tensor):
    need = C.c_int(1)
    rows, cols = x32.shape
    N.check(N.load().n2v2r_planes_needed(N.stream_ptr(), N.ptr(x32), rows, cols, x32.stride(0), C.byref(need)))
    return need.value


def transform(x: torch.Tensor, absolute=False, threshold=None, cut=None, binarize=False,
              want_f32=False, n_planes=0, want_colsum=False):
    """K1b fused load pass.  Returns (out_f32 | None, planes | None, ld_p, colsum | None)."""
    lib = N.load()
    rows, cols = x.shape
    out = torch.empty((rows, cols), dtype=torch.float32, device=x.device) if want_f32 else None
    ld_p = (cols + 63) // 64 * 64
    planes = torch.empty((n_planes, rows, ld_p), dtype=torch.bfloat16, device=x.device) if n_planes else None
    colsum = torch.zeros(cols, dtype=torch.float64, device=x.device) if want_colsum else None
    N.check(lib.n2v2r_transform(N.stream_ptr(), N.ptr(x), _dtype_code(x), rows, cols, x.stride(0),
                                int(bool(absolute)), int(threshold is not None), float(threshold or 0.0),
                                int(cut is not None), float(cut if cut is not None else 0.0), int(bool(binarize)),
                                N.ptr(out), cols, N.ptr(planes), n_planes, ld_p, N.ptr(colsum)))
    return out, planes, ld_p, colsum


def project(W: torch.Tensor, on):
    """K2: W^T W ('columns') or W W^T ('rows') on the tensor cores -> fp32 device matrix."""
    lib = N.load()
    which = on.casefold()                       # AttributeError on None, like the reference
    if which not in ("columns", "rows"):
        raise ValueError("Unknown projection type, options are columns or rows")
    rows, cols = W.shape
    assert rows != cols, "Graph is not bipartite"
    on_rows = int(which == "rows")
    m = rows if on_rows else cols
    ws = torch.empty(lib.n2v2r_project_workspace_bytes(rows, cols, on_rows), dtype=torch.uint8, device=W.device)
    out = torch.empty((m, m), dtype=torch.float32, device=W.device)
    N.check(lib.n2v2r_project(N.stream_ptr(), N.ptr(W), _dtype_code(W), rows, cols, W.stride(0), on_rows,
                              N.ptr(out), m, N.ptr(ws), ws.numel()))
    return out


E4M3_AUTO = True                # storage="auto": large binarised layers are held as one e4m3 plane (C2: 9-16 % faster steps than
                                # the bf16 plane, profiles/README.md).  Seed batching needs bf16 planes: pass storage="planes"
def make_slab_layer(x32: torch.Tensor, row_offset: int, rows: int, n_planes=None):
    """Row slab of a layer for sharding ONE layer over several GPUs (fewer layers than GPUs): x32 is the (n, n)
    fp32 device matrix (or just its rows [row_offset, row_offset + rows)); the slab and its transpose are stored as bf16
    planes.  The solver (device.uase with a group communicator) all-reduces Z_t over the ranks that hold the other slabs."""
    n = x32.shape[1]
    block = x32 if x32.shape[0] == rows else x32[row_offset:row_offset + rows]
    if block.shape != (rows, n):
        raise ValueError("make_slab_layer: rows outside the matrix")
    block = block.contiguous()
    P = n_planes or planes_needed(block)
    _, planes, ld_p, cs = transform(block, n_planes=P, want_colsum=True)                # (P, rows, ld_n)
    ld_t = (rows + 63) // 64 * 64
    planes_t = torch.zeros((P, n, ld_t), dtype=torch.bfloat16, device=block.device)
    planes_t[:, :, :rows] = planes[:, :, :n].transpose(1, 2)                            # build-time plumbing, not on the hot path
    return DeviceLayer(planes=planes, n=n, ld=ld_p, n_planes=P, symmetric=False, planes_t=planes_t, colsum=cs,
                       row_offset=int(row_offset), rows=int(rows), ld_t=ld_t)


TILES_MAX_DENSITY = 0.08        # layers sparser than this are stored as compressed tiles (storage="auto")
TILES_MIN_N = 4096              # ... when they are large enough for the streaming product to matter


def make_tiles_layer(x32: torch.Tensor, colsum=None, max_density=None, tile_rows=None):
    """fp32 (n, n) device matrix -> DeviceLayer(storage="tiles"), or (None, planes, symmetric) when the matrix is denser
    than `max_density` (the caller then builds planes; the count pass has already told it how many and the symmetry)."""
    lib = N.load()
    n = x32.shape[0]
    assert x32.dim() == 2 and x32.shape[0] == x32.shape[1] and x32.dtype == torch.float32, "layer must be a square float32 matrix"
    if x32.stride(1) != 1:
        x32 = x32.contiguous()
    dev = x32.device
    n_kb, n_rt128 = (n + 63) // 64, (n + 127) // 128
    counts = torch.empty((2, n_rt128 * n_kb), dtype=torch.int32, device=dev)
    planes, sym = C.c_int(1), C.c_int(1)
    N.check(lib.n2v2r_tiles_count(N.stream_ptr(), N.ptr(x32), n, x32.stride(0), N.ptr(counts[0]), N.ptr(counts[1]),
                                  C.byref(planes), C.byref(sym)))
    P = planes.value
    tile_rows = int(tile_rows) if tile_rows else (256 if P <= 2 else 128)
    if tile_rows not in (128, 256) or (tile_rows == 256 and P == 3):
        raise ValueError("tile_rows must be 128 or 256 (128 for three-plane layers)")
    n_tiles = ((n + tile_rows - 1) // tile_rows) * n_kb

    def build(which, transpose):
        tp = torch.empty(n_tiles + 1, dtype=torch.int32, device=dev)
        total = C.c_int64(0)
        N.check(lib.n2v2r_tiles_scan(N.stream_ptr(), N.ptr(counts[which]), n, tile_rows, N.ptr(tp), C.byref(total)))
        if max_density is not None and total.value > max_density * n * n:
            return None, None, 0
        ent = torch.empty((max(total.value, 2), 2), dtype=torch.int32, device=dev)
        N.check(lib.n2v2r_tiles_fill(N.stream_ptr(), N.ptr(x32), n, x32.stride(0), int(transpose), tile_rows, N.ptr(tp), N.ptr(ent)))
        return tp, ent, total.value

    tp, ent, cnt = build(0, False)
    if tp is None:
        return None, P, bool(sym.value)
    tp_t = ent_t = None
    cnt_t = 0
    if not sym.value:
        tp_t, ent_t, cnt_t = build(1, True)
    return DeviceLayer(planes=None, n=n, ld=(n + 63) // 64 * 64, n_planes=P, symmetric=bool(sym.value), colsum=colsum,
                       storage="tiles", tile_rows=tile_rows, tile_ptr=tp, entries=ent, tile_ptr_t=tp_t, entries_t=ent_t,
                       n_entries=cnt, n_entries_t=cnt_t), P, bool(sym.value)


ROWS_MAX_DENSITY = 0.045        # layers sparser than this are stored as row lists (storage="auto"): the fp32 FMA-pipe product costs
                                # about 0.26 ms per percent of density at n = 50 000 against 0.96 ms for the tiles at any density
ROWS_PANEL_COLS = 256           # columns per panel when the solver's block width is not known at build time: the X panel (256 x l
                                # fp32) and the block's list make one stage in shared memory, and two stages fit for every l <= 80
ROWS_MAX_LIST_BYTES = 36 * 1024  # longest (block, panel) list a layer may have then: two stages must fit beside two X panels
ROWS_STAGE_BYTES = 113 * 1024   # half of the shared memory the product kernel can use: one stage = X panel + longest list
ROWS_GROUPS = 176               # groups of 4 lanes per CTA of the product kernel (22 warps): rows per block <= 3 x this


_SM_COUNT = {}


def sm_count():
    dev = torch.cuda.current_device()
    if dev not in _SM_COUNT:
        sms, cc, mem = C.c_int(0), C.c_int(0), C.c_size_t(0)
        N.check(N.load().n2v2r_device_info(dev, C.byref(sms), C.byref(cc), C.byref(mem)))
        _SM_COUNT[dev] = sms.value
    return _SM_COUNT[dev]


def rows_panel_cols(block_width):
    """Columns per panel of a row-list layer built for a solver block of `block_width` columns: the widest panel (multiple of
    64, at most 512) whose stage -- X panel + a list of about 80 bytes per column at C3's 2.6 % -- fits twice in shared memory.
    Wider panels mean longer row segments per panel, i.e. fewer idle lane-group steps: a C3 pass at l = 32 takes 0.463 / 0.441 /
    0.427 / 0.416 / 0.407 ms with 256 / 320 / 384 / 448 / 512 columns.  A layer built for a block width also works with wider
    blocks as long as ONE stage fits (the kernel then runs without overlap of copy and compute)."""
    if not block_width:
        return ROWS_PANEL_COLS
    w = ROWS_STAGE_BYTES // (4 * int(block_width) + 96)
    return int(max(128, min(512, w // 64 * 64)))


def rows_per_block(n):
    """Rows per block of a row-list layer: one block per SM when the groups can own that many rows (<= 3 each), else the
    smallest split into equal waves."""
    sms = sm_count()
    waves = 1
    while (n + sms * waves - 1) // (sms * waves) > 3 * ROWS_GROUPS:
        waves += 1
    return max(1, (n + sms * waves - 1) // (sms * waves))


def make_rows_layer(x32: torch.Tensor, colsum=None, max_density=None, rows_block=None, panel_cols=None, max_list_bytes=ROWS_MAX_LIST_BYTES,
                    block_width=None):
    """fp32 (n, n) device matrix -> DeviceLayer(storage="rows"), or None when the matrix is denser than `max_density` or one
    of its (block, panel) lists is longer than `max_list_bytes` (a hub row block: the caller picks another storage).
    block_width: the solver block the layer will be multiplied with, when known: sets the panel width (rows_panel_cols) and
    the list bound (what is left of a stage beside the X panel)."""
    lib = N.load()
    n = x32.shape[0]
    assert x32.dim() == 2 and x32.shape[0] == x32.shape[1] and x32.dtype == torch.float32, "layer must be a square float32 matrix"
    if x32.stride(1) != 1:
        x32 = x32.contiguous()
    dev = x32.device
    rb = int(rows_block) if rows_block else rows_per_block(n)
    W = int(panel_cols) if panel_cols else rows_panel_cols(block_width)
    if block_width and not panel_cols and max_list_bytes == ROWS_MAX_LIST_BYTES:
        max_list_bytes = ROWS_STAGE_BYTES - W * 4 * int(block_width)
    n_lists = ((n + rb - 1) // rb) * ((n + W - 1) // W)

    def build(src):
        counts = torch.empty(lib.n2v2r_rows_counts_bytes(n, rb, W), dtype=torch.uint8, device=dev)
        lp = torch.empty(n_lists + 1, dtype=torch.int32, device=dev)
        units, cap, nnz = C.c_int64(0), C.c_int32(0), C.c_int64(0)
        N.check(lib.n2v2r_rows_count(N.stream_ptr(), N.ptr(src), n, src.stride(0), rb, W, N.ptr(counts), N.ptr(lp),
                                     C.byref(units), C.byref(cap), C.byref(nnz)))
        if (max_density is not None and nnz.value > max_density * n * n) or (max_list_bytes and cap.value > max_list_bytes):
            return None
        lists = torch.empty((max(units.value, 1), 4), dtype=torch.int32, device=dev)
        N.check(lib.n2v2r_rows_fill(N.stream_ptr(), N.ptr(src), n, src.stride(0), rb, W, N.ptr(counts), N.ptr(lp), N.ptr(lists)))
        return lp, lists, units.value, cap.value, nnz.value

    a = build(x32)
    if a is None:
        return None
    sym = bool(torch.equal(x32, x32.t()))
    at = None
    if not sym:
        at = build(x32.t().contiguous())
        if at is None:
            return None
    if colsum is None:
        colsum = x32.sum(dim=0, dtype=torch.float64)
    return DeviceLayer(planes=None, n=n, ld=(n + 63) // 64 * 64, n_planes=0, symmetric=sym, colsum=colsum, storage="rows",
                       tile_rows=rb, panel_cols=W, tile_ptr=a[0], entries=a[1], n_units=a[2], list_cap=a[3], n_entries=a[4],
                       tile_ptr_t=None if at is None else at[0], entries_t=None if at is None else at[1],
                       n_units_t=0 if at is None else at[2], list_cap_t=0 if at is None else at[3], n_entries_t=0 if at is None else at[4])


def make_layer(x32: torch.Tensor, n_planes=None, colsum=None, check_symmetry=True, storage="auto", block_width=None):
    """fp32 (n, n) device matrix -> DeviceLayer.  storage "auto" for n >= TILES_MIN_N: row lists when the density is <=
    ROWS_MAX_DENSITY (and no list is too long for a stage), compressed tiles when it is <= TILES_MAX_DENSITY, one e4m3 plane
    for binarised layers, else bf16 planes + symmetry flag / transposed copy.  "rows" / "tiles" / "planes" / "e4m3" force one."""
    n = x32.shape[0]
    assert x32.shape[0] == x32.shape[1], "layer must be square"
    if storage not in ("auto", "planes", "tiles", "e4m3", "rows"):
        raise ValueError(f"unknown layer storage {storage!r}")
    if storage == "e4m3":
        return make_layer(x32, n_planes=1, colsum=colsum, check_symmetry=check_symmetry, storage="planes").to_e4m3()
    if storage == "rows" or (storage == "auto" and n >= TILES_MIN_N and n_planes is None):
        layer = make_rows_layer(x32, colsum=colsum, max_density=None if storage == "rows" else ROWS_MAX_DENSITY,
                                max_list_bytes=None if storage == "rows" else ROWS_MAX_LIST_BYTES, block_width=block_width)
        if layer is not None:
            return layer
        if storage == "rows":
            raise ValueError("layer cannot be stored as row lists")
    known_sym = None
    if storage == "tiles" or (storage == "auto" and n >= TILES_MIN_N and n_planes is None):
        layer, P, known_sym = make_tiles_layer(x32, colsum=colsum, max_density=None if storage == "tiles" else TILES_MAX_DENSITY)
        if layer is not None:
            if layer.colsum is None:
                layer.colsum = x32.sum(dim=0, dtype=torch.float64)
            return layer
        n_planes = n_planes or P
    explicit_planes = n_planes is not None and known_sym is None
    P = n_planes or planes_needed(x32)
    _, planes, ld_p, cs = transform(x32, n_planes=P, want_colsum=colsum is None)
    layer = finish_layer(planes, n, ld_p, P, cs if colsum is None else colsum, check_symmetry, known_symmetric=known_sym)
    if storage == "auto" and E4M3_AUTO and P == 1 and n >= TILES_MIN_N and not explicit_planes:
        try:
            layer = layer.to_e4m3()            # one-plane layers whose values e4m3 holds exactly (binarised): half the bytes
        except ValueError:
            pass
    return layer


PACK_MIN_ENTRIES = 4_000_000          # below this a layer crosses PCIe dense (the packed pipeline has fixed costs)
PACK_CHUNK_ENTRIES = 300_000_000      # rows are packed and copied in chunks of about this many entries
PACK_MAX_DENSITY = 0.3                # staging capacity for the non-zero values, as a fraction of a chunk's entries
PACK_SLOTS = 3                        # pinned staging slots: packed chunks in flight behind a dense chunk on the copy engine
PACK_DENSE_EVERY = 0                  # N > 1 (pinned sources): every N-th row chunk crosses PCIe dense on its own stream while the
                                      # host cores pack the others.  Measured at C3 (profiles/README.md): with N = 4 the DMA reads
                                      # compete with the packing threads for host memory bandwidth and the step gets SLOWER
                                      # (transfer 0.73 s against 0.57 s), so every chunk is packed (0) by default
_pack_staging = {}                    # (chunk_rows, cols) -> two pinned staging slots, kept across calls (pinning is slow)


def host_threads():
    """Host threads for the packing: the cores this process may use, shared between the ranks of one node."""
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    try:
        local_world = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    except ValueError:
        local_world = 1
    return max(1, cores // local_world)


def _staging(chunk_rows, cols):
    words = (cols + 31) // 32
    cap = max(8, int(chunk_rows * cols * PACK_MAX_DENSITY))
    key = (chunk_rows, cols, cap)
    if key not in _pack_staging:
        _pack_staging.clear()                                      # one shape at a time stays pinned
        _pack_staging[key] = [dict(bitmap=torch.empty((chunk_rows, words), dtype=torch.int32, pin_memory=True),
                                   values=torch.empty(cap, dtype=torch.float32, pin_memory=True),
                                   row_start=torch.empty(chunk_rows, dtype=torch.int64, pin_memory=True)) for _ in range(PACK_SLOTS)]
    return _pack_staging[key]


def layers_from_host(arrays, n_planes=None, transfer="auto", storage="auto", block_width=None):
    """float32 (n, n) host matrices (what the reference's DataLoader hands to N2V2R) -> DeviceLayers.

    transfer = "dense": double-buffered H2D of the dense matrices (the copy of layer t+1 runs on a copy stream while
    the load pass of layer t runs on the current stream).  Pinned sources copy asynchronously; pageable ones block.
    transfer = "packed": the layers cross PCIe as bitmap + non-zero values (K0c): a producer thread packs chunks of
    rows with every host core (n2v2r_host_pack_nonzeros), copies them and expands them on the copy stream into one of
    two dense device buffers while the current stream runs the load pass of the previous layer.  The reference does
    the same dense -> sparse conversion on the host (csc_matrix, model.py:51-53).  A chunk denser than
    PACK_MAX_DENSITY is copied dense, and so is everything after it.
    transfer = "auto": packed for layers of at least PACK_MIN_ENTRIES entries, dense below."""
    srcs = [torch.from_numpy(np.ascontiguousarray(a)) for a in arrays]
    if not srcs:
        return []
    if any(x.dtype != torch.float32 or x.shape != srcs[0].shape or x.dim() != 2 for x in srcs):
        raise TypeError("layers_from_host expects float32 matrices of one shape")
    if transfer not in ("auto", "dense", "packed"):
        raise ValueError(f"unknown transfer mode {transfer!r}")
    if transfer == "auto":
        transfer = "packed" if srcs[0].numel() >= PACK_MIN_ENTRIES else "dense"
    if transfer == "packed":
        rows, cols = srcs[0].shape
        chunk_rows = max(1, min(rows, PACK_CHUNK_ENTRIES // max(cols, 1)))
        try:
            slots = _staging(chunk_rows, cols)
        except RuntimeError:                               # no lockable host memory for the staging slots: dense copies
            slots = None
        if slots is not None:
            return _layers_from_host_packed(srcs, n_planes, chunk_rows, slots, storage, block_width)
    return _layers_from_host_dense(srcs, n_planes, storage, block_width)


def _layers_from_host_packed(srcs, n_planes, chunk_rows, slots, storage="auto", block_width=None):
    import queue
    import threading
    dev = _dev()
    lib = N.load()
    rows, cols = srcs[0].shape
    words = (cols + 31) // 32
    cur = torch.cuda.current_stream(dev)
    copy = torch.cuda.Stream(dev)
    copy2 = torch.cuda.Stream(dev)                        # dense chunks of the hybrid transfer
    hybrid = PACK_DENSE_EVERY > 1 and all(x.is_pinned() for x in srcs) and -(-rows // chunk_rows) >= PACK_DENSE_EVERY
    dense = [torch.empty((rows, cols), dtype=torch.float32, device=dev) for _ in range(min(2, len(srcs)))]
    twins = [{k: torch.empty_like(v, device=dev) for k, v in slot.items()} for slot in slots]
    copy.wait_stream(cur)                                 # the buffers were allocated on the current stream
    copy2.wait_stream(cur)
    ready, consumed = queue.Queue(), queue.Queue()
    threads = host_threads()
    stats = dict(packed_chunks=0, dense_chunks=0, h2d_bytes=0, pack_s=0.0, producer_s=0.0, total_s=0.0, threads=threads)
    stats["hybrid"] = bool(hybrid)
    t_begin = time.perf_counter()

    def produce():
        try:
            torch.cuda.set_device(dev)
            sent = []                                     # per chunk: event after its H2D (the staging slot is free again)
            dense_mode = False
            with torch.cuda.stream(copy):
                for t, src in enumerate(srcs):
                    if t >= 2:
                        done = consumed.get()             # the load pass of layer t-2 has read this dense buffer
                        if done is None:                  # the consumer failed
                            return
                        copy.wait_event(done)
                        copy2.wait_event(done)
                    out = dense[t % 2]
                    for ci, r0 in enumerate(range(0, rows, chunk_rows)):
                        r1 = min(rows, r0 + chunk_rows)
                        nnz = C.c_int64(-1)
                        if hybrid and not dense_mode and ci % PACK_DENSE_EVERY == PACK_DENSE_EVERY - 1:
                            # this chunk crosses PCIe as it is, queued on its own stream in chunk order: the copy engine works
                            # on it while the host cores pack the next chunks (their copies queue up behind it)
                            with torch.cuda.stream(copy2):
                                out[r0:r1].copy_(src[r0:r1], non_blocking=True)
                            stats["dense_chunks"] += 1
                            stats["h2d_bytes"] += (r1 - r0) * cols * 4
                            continue
                        if not dense_mode:
                            c = len(sent)
                            slot, twin = slots[c % PACK_SLOTS], twins[c % PACK_SLOTS]
                            if c >= PACK_SLOTS:
                                sent[c - PACK_SLOTS].synchronize()
                            t_pack = time.perf_counter()
                            N.check(lib.n2v2r_host_pack_nonzeros(C.c_void_p(src[r0:r1].data_ptr()), r1 - r0, cols, cols,
                                                                 N.ptr(slot["bitmap"]), N.ptr(slot["values"]),
                                                                 slot["values"].numel() - 8, N.ptr(slot["row_start"]),
                                                                 threads, C.byref(nnz)))
                            stats["pack_s"] += time.perf_counter() - t_pack
                        if nnz.value < 0:                 # too dense (or dense mode): this chunk crosses PCIe as it is
                            dense_mode = True
                            out[r0:r1].copy_(src[r0:r1], non_blocking=True)
                            stats["dense_chunks"] += 1
                            stats["h2d_bytes"] += (r1 - r0) * cols * 4
                            continue
                        k = nnz.value
                        twin["bitmap"][: r1 - r0].copy_(slot["bitmap"][: r1 - r0], non_blocking=True)
                        twin["row_start"][: r1 - r0].copy_(slot["row_start"][: r1 - r0], non_blocking=True)
                        twin["values"][:k].copy_(slot["values"][:k], non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(copy)
                        sent.append(ev)
                        N.check(lib.n2v2r_unpack_nonzeros(C.c_void_p(copy.cuda_stream), N.ptr(twin["bitmap"]), N.ptr(twin["values"]),
                                                          N.ptr(twin["row_start"]), r1 - r0, cols, N.ptr(out[r0:r1]), cols))
                        stats["packed_chunks"] += 1
                        stats["h2d_bytes"] += (r1 - r0) * (words * 4 + 8) + k * 4
                    copy.wait_stream(copy2)
                    ev = torch.cuda.Event()
                    ev.record(copy)
                    ready.put(ev)
            stats["producer_s"] = time.perf_counter() - t_begin
        except BaseException as ex:  # noqa: BLE001  (handed to the consumer)
            ready.put(ex)

    worker = threading.Thread(target=produce, name="n2v2r-pack", daemon=True)
    worker.start()
    layers = []
    try:
        for t in range(len(srcs)):
            ev = ready.get()
            if isinstance(ev, BaseException):
                raise ev
            cur.wait_event(ev)
            layers.append(make_layer(dense[t % 2], n_planes=n_planes, storage=storage, block_width=block_width))
            done = torch.cuda.Event()
            done.record(cur)
            consumed.put(done)
    except BaseException:
        consumed.put(None)
        consumed.put(None)
        raise
    finally:
        worker.join()
        cur.wait_stream(copy2)
        cur.wait_stream(copy)          # also on the error path: the buffers go back to the allocator only after the copy stream
    stats["total_s"] = time.perf_counter() - t_begin
    layers_from_host.last_stats = stats
    return layers


def _layers_from_host_dense(srcs, n_planes, storage="auto", block_width=None):
    dev = _dev()
    t_begin = time.perf_counter()
    stats = dict(packed_chunks=0, dense_chunks=len(srcs), h2d_bytes=sum(x.numel() * 4 for x in srcs), total_s=0.0)
    layers_from_host.last_stats = stats
    cur = torch.cuda.current_stream(dev)
    copy = torch.cuda.Stream(dev)
    bufs = [torch.empty(srcs[0].shape, dtype=torch.float32, device=dev) for _ in range(min(2, len(srcs)))]
    copied = [torch.cuda.Event() for _ in srcs]
    built = [torch.cuda.Event() for _ in srcs]

    def issue(t):
        with torch.cuda.stream(copy):
            if t >= 2:
                copy.wait_event(built[t - 2])          # the buffer's previous layer has been consumed
            else:
                copy.wait_stream(cur)                  # buffers were allocated on the current stream
            bufs[t % 2].copy_(srcs[t], non_blocking=True)
            copied[t].record(copy)

    issue(0)
    layers = []
    for t in range(len(srcs)):
        if t + 1 < len(srcs):
            issue(t + 1)
        cur.wait_event(copied[t])
        layers.append(make_layer(bufs[t % 2], n_planes=n_planes, storage=storage, block_width=block_width))
        built[t].record(cur)
    cur.wait_stream(copy)
    stats["total_s"] = time.perf_counter() - t_begin        # host time: the last load pass may still be running
    return layers


def finish_layer(planes, n, ld_p, P, colsum, check_symmetry=True, known_symmetric=None):
    lib = N.load()
    sym = C.c_int(1)
    planes_t = None
    if check_symmetry:
        if known_symmetric is not None:
            sym = C.c_int(int(known_symmetric))
        else:
            N.check(lib.n2v2r_is_symmetric(N.stream_ptr(), N.ptr(planes), P, n, ld_p, C.byref(sym)))
        if not sym.value:
            planes_t = torch.empty_like(planes)
            N.check(lib.n2v2r_transpose_planes(N.stream_ptr(), N.ptr(planes), P, n, ld_p, N.ptr(planes_t)))
    return DeviceLayer(planes=planes, n=n, ld=ld_p, n_planes=P, symmetric=bool(sym.value), planes_t=planes_t, colsum=colsum)


def preprocess_to_layer(x: torch.Tensor, threshold=None, top_percent_keep=100, binarize=False, absolute=False,
                        project_unipartite_on="columns", storage_planes=None, want_f32=False, storage="auto", block_width=None):
    """network_transform (preprocessing_utils.py:116-141) on the device.

    Returns (DeviceLayer, out_f32 | None).  Order of operations as the reference: abs ->
    threshold -> projection when non-square -> percentile cut over the nonzeros -> binarize
    -> fp32.  The select and apply kernels share one element transform, so the cut is taken
    on exactly the values it is applied to."""
    rows, cols = x.shape
    projected = rows != cols
    if projected:
        if absolute or threshold is not None:
            x, _, _, _ = transform(x, absolute=absolute, threshold=threshold, want_f32=True)
        x = project(x, project_unipartite_on)
        absolute, threshold = False, None
    # the reference takes the percentile even for keep == 100 (:134); q = 0 is the minimum of the nonzeros, below which
    # nothing lies, so the cut is a no-op there; for a projected layer (a Gram matrix, never all-zero unless W is) the
    # select passes are skipped, square inputs keep them so that an all-zero graph raises like np.percentile does
    if float(top_percent_keep) == 100.0 and projected:
        cut = None
    else:
        cut = percentile_cut(x, 100 - top_percent_keep, absolute, threshold)
    n = x.shape[0]
    check_sym = not projected            # K2 mirrors the upper triangle: W^T W is exactly symmetric, and K1b is elementwise
    if storage == "e4m3" and not binarize:
        raise ValueError('layer storage "e4m3" needs binarised layers (values exactly representable in e4m3)')
    if binarize and storage not in ("tiles", "rows"):
        P = 1
    elif storage_planes and storage not in ("tiles", "rows"):
        P = int(storage_planes)
    else:
        P = 0
    if P == 0:
        # exact storage: materialise fp32 once, count the planes it needs, then split
        if cut is None and not binarize and not absolute and threshold is None:
            layer = make_layer(x, check_symmetry=check_sym, storage=storage, block_width=block_width)   # identity transform: store the projected matrix directly
            return layer, (x if want_f32 else None)
        out, _, _, cs = transform(x, absolute, threshold, cut, binarize, want_f32=True, want_colsum=True)
        layer = make_layer(out, colsum=cs, check_symmetry=check_sym, storage=storage, block_width=block_width)
        return layer, (out if want_f32 else None)
    out, planes, ld_p, cs = transform(x, absolute, threshold, cut, binarize, want_f32=want_f32, n_planes=P,
                                      want_colsum=True)
    layer = finish_layer(planes, n, ld_p, P, cs, check_sym)
    if storage == "e4m3" or (storage == "auto" and binarize and E4M3_AUTO and n >= TILES_MIN_N):
        layer = layer.to_e4m3()          # {0, 1} is exact in e4m3: half the bytes of the bf16 plane
    return layer, out


CONVERGED_STATUS = {0: "max_iters", 1: "converged", 2: "stagnated"}


def uase(layers, k, l=64, tol=1e-5, max_iters=300, min_iters=2, seed=42, check_dims=None, algo=N.ALGO_TCGEN05,
         k_chunk=0, allreduce=None, T_global=None, layer_offset=0, want_left=False, cheb_degree=0, group=None,
         floor_mult=0.0, rel_floor=0.0, first_degree=0, no_early_lock=False, trace=False, slab_group=None):
    """K3-K8: returns (embeddings (T_local, n, k) fp32 device, sigma (k,) float64, left | None, stats dict).

    stats["converged"] is the solver's tri-state (0 = max_iters reached, 1 = angle bound <= tol at every checked d or every
    wanted residual at its target, 2 = stopped because the residuals stagnated above the target), stats["converged_status"]
    its name, stats["bound_met"] whether the Davis-Kahan bound met `tol`, stats["angle_bounds"] the bound per checked d."""
    lib = N.load()
    dev = _dev()
    T = len(layers)
    n = layers[0].n
    dims = np.asarray(sorted(set(int(d) for d in (check_dims or [k]) if 1 <= int(d) <= k)), dtype=np.int32)
    if len(dims) > N.MAX_CHECK:
        dims = dims[-N.MAX_CHECK:]
    p = N.UaseParams(n=n, T_local=T, T_global=T_global or T, layer_offset=layer_offset, k=k, l=l,
                     max_iters=max_iters, min_iters=min_iters, cheb_degree=cheb_degree, algo=algo, k_chunk=k_chunk,
                     n_check=len(dims), tol=tol, seed=seed,
                     h_check_dims=dims.ctypes.data_as(C.POINTER(C.c_int32)),
                     floor_mult=float(floor_mult), rel_floor=float(rel_floor), first_degree=int(first_degree),
                     no_early_lock=int(bool(no_early_lock)), trace=int(bool(trace)), comm_mode=0)
    ws_bytes = lib.n2v2r_uase_workspace_bytes(C.byref(p))
    if ws_bytes == 0:
        raise N.NativeError(N.EINVAL, lib.n2v2r_last_error().decode())
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    emb = torch.empty((T, n, k), dtype=torch.float32, device=dev)
    left = torch.empty((n, k), dtype=torch.float32, device=dev) if want_left else None
    sigma = np.zeros(k, dtype=np.float64)
    stats = N.UaseStats()
    arr = (N.Layer * T)(*[L.c_struct() for L in layers])
    state = None
    user = None
    slabs = any(L.rows > 0 for L in layers)
    if slabs:                                    # row slabs: Y over all ranks, Z_t over the ranks of this layer group
        from . import dist
        comm = dist.native_comm(group)
        if comm is None:
            raise RuntimeError("row-slab layers need torch.distributed on NCCL (the library issues the all-reduces itself)")
        p.comm_mode = 1
        user = comm
        p.group_comm = dist.native_comm(slab_group) if slab_group is not None else comm
        cb = C.cast(None, N.ALLREDUCE_FN)
    elif allreduce is not None:
        cb = N.ALLREDUCE_FN(allreduce)
    elif (T_global or T) > T:                    # layer-sharded: all-reduce Y over the ranks (dist.py)
        from . import dist
        comm = dist.native_comm(group)
        if comm is not None:                     # ncclAllReduce issued by the library itself, no Python in the solver loop
            p.comm_mode = 1
            user = comm
            cb = C.cast(None, N.ALLREDUCE_FN)
        else:
            cb, state = dist.make_allreduce_callback(ws, group)
    else:
        cb = C.cast(None, N.ALLREDUCE_FN)
    rc = lib.n2v2r_uase(N.stream_ptr(), C.byref(p), arr, N.ptr(ws), ws_bytes, cb, user, N.ptr(emb), N.ptr(left),
                        sigma.ctypes.data_as(C.POINTER(C.c_double)), C.byref(stats))
    if state is not None and state["error"] is not None:
        raise state["error"]
    N.check(rc, allow=(N.ENOTCONV,))
    info = dict(allreduce_calls=(stats.iterations if p.comm_mode == 1 else 0) if state is None else state["calls"],
                allreduce_bytes=(stats.iterations * n * l * 4 if p.comm_mode == 1 else 0) if state is None else state["bytes"],
                allreduce_path="nccl-native" if p.comm_mode == 1 else ("callback" if state is not None else "none"),
                iterations=stats.iterations, layer_passes=stats.layer_passes, converged=int(stats.converged),
                converged_status=CONVERGED_STATUS.get(int(stats.converged), "?"), bound_met=bool(stats.bound_met),
                kernel_launches=stats.kernel_launches, outer_iterations=stats.outer_iterations,
                locked=stats.locked, stream_ms=stats.stream_ms, mma_ms=stats.mma_ms, mma_launches=stats.mma_launches, total_ms=stats.total_ms,
                stream_bytes=stats.stream_bytes, angle_bound=stats.max_angle_bound,
                angle_bounds={int(d): float(stats.angle_bounds[i]) for i, d in enumerate(dims)},
                resid=np.asarray(stats.resid[:l]), ritz=np.asarray(stats.ritz[:l]), status=rc, l=l)
    return emb, sigma, left, info


def uase_batched(layers, k, seeds, l=64, tol=1e-5, max_iters=300, min_iters=2, check_dims=None, k_chunk=0, cheb_degree=0):
    """Seed batching (stability runs): one UASE solve per seed on the same resident layers, run CONCURRENTLY -- a host thread
    and a CUDA stream per solve -- with every product M X served for all waiting solves by one pass over each layer per up
    to three operands (n2v2r_batch_*, csrc/batch.cu).  Each solve is bit-identical to device.uase(layers, k, seed=...).
    Returns ([(emb, sigma, info)] in the order of `seeds`, batch stats dict)."""
    import threading
    lib = N.load()
    dev = _dev()
    seeds = [int(s) for s in seeds]
    S = len(seeds)
    if S == 0:
        return [], {}
    T, n = len(layers), layers[0].n
    dims = np.asarray(sorted(set(int(d) for d in (check_dims or [k]) if 1 <= int(d) <= k)), dtype=np.int32)[-N.MAX_CHECK:]
    arr = (N.Layer * T)(*[L.c_struct() for L in layers])

    def params(seed=0, batch=None, slot=0):
        return N.UaseParams(n=n, T_local=T, T_global=T, layer_offset=0, k=k, l=l, max_iters=max_iters, min_iters=min_iters,
                            cheb_degree=cheb_degree, algo=N.ALGO_TCGEN05, k_chunk=k_chunk, n_check=len(dims), tol=tol, seed=seed,
                            h_check_dims=dims.ctypes.data_as(C.POINTER(C.c_int32)), batch=batch, batch_slot=slot)

    p0 = params()
    ws_bytes = lib.n2v2r_uase_workspace_bytes(C.byref(p0))
    if ws_bytes == 0:
        raise N.NativeError(N.EINVAL, lib.n2v2r_last_error().decode())
    bws = torch.empty(lib.n2v2r_batch_workspace_bytes(C.byref(p0), S), dtype=torch.uint8, device=dev)
    handle = C.c_void_p()
    N.check(lib.n2v2r_batch_create(C.byref(p0), arr, S, N.ptr(bws), bws.numel(), C.byref(handle)))
    main = torch.cuda.current_stream(dev)
    # every buffer is allocated on the current stream before the solves start; the solve streams wait for it and are waited for
    wss = [torch.empty(ws_bytes, dtype=torch.uint8, device=dev) for _ in range(S)]
    embs = [torch.empty((T, n, k), dtype=torch.float32, device=dev) for _ in range(S)]
    sigmas = [np.zeros(k, dtype=np.float64) for _ in range(S)]
    stats = [N.UaseStats() for _ in range(S)]
    streams = [torch.cuda.Stream(dev) for _ in range(S)]
    rcs, errs = [None] * S, [None] * S
    for st in streams:
        st.wait_stream(main)

    def run(i):
        try:
            torch.cuda.set_device(dev)
            p = params(seeds[i], handle, i)
            rcs[i] = lib.n2v2r_uase(C.c_void_p(streams[i].cuda_stream), C.byref(p), arr, N.ptr(wss[i]), ws_bytes,
                                    C.cast(None, N.ALLREDUCE_FN), None, N.ptr(embs[i]), None,
                                    sigmas[i].ctypes.data_as(C.POINTER(C.c_double)), C.byref(stats[i]))
            if rcs[i] not in (N.OK, N.ENOTCONV):
                errs[i] = N.NativeError(rcs[i], lib.n2v2r_last_error().decode("utf-8", "replace"))
        except BaseException as ex:  # noqa: BLE001
            errs[i] = ex

    threads = [threading.Thread(target=run, args=(i,), name=f"n2v2r-seed-{seeds[i]}") for i in range(S)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for st in streams:
        main.wait_stream(st)
    b4 = (C.c_double * 4)()
    brc = lib.n2v2r_batch_destroy(handle, b4)
    for e in errs:
        if e is not None:
            raise e
    N.check(brc)
    out = []
    for i in range(S):
        st_ = stats[i]
        info = dict(seed=seeds[i], iterations=st_.iterations, layer_passes=st_.layer_passes, converged=int(st_.converged),
                    converged_status=CONVERGED_STATUS.get(int(st_.converged), "?"), bound_met=bool(st_.bound_met),
                    kernel_launches=st_.kernel_launches, outer_iterations=st_.outer_iterations, locked=st_.locked,
                    stream_ms=st_.stream_ms, mma_ms=b4[2] / S, mma_launches=b4[3] / S, total_ms=st_.total_ms, stream_bytes=st_.stream_bytes,
                    angle_bound=st_.max_angle_bound, angle_bounds={int(d): float(st_.angle_bounds[j]) for j, d in enumerate(dims)},
                    resid=np.asarray(st_.resid[:l]), ritz=np.asarray(st_.ritz[:l]), status=rcs[i], l=l,
                    allreduce_calls=0, allreduce_bytes=0, allreduce_path="none")
        out.append((embs[i], sigmas[i], info))
    batch = dict(products=int(b4[0]), layer_streams=int(b4[1]), mma_ms=float(b4[2]), mma_launches=int(b4[3]),
                 unbatched_layer_streams=int(sum(s_.layer_passes for s_ in stats)))
    return out, batch


def subspace_sin_angles(emb_a: torch.Tensor, emb_b: torch.Tensor, dims):
    """sin of the largest principal angle between span(V_a[:, :d]) and span(V_b[:, :d]) for every d in dims, V = the
    stacked (T n x k) embeddings of two solves.  A CHECKER (bench / tests), not on the hot path: three k x k Gram matrices
    accumulated in float64 on the device (torch), the rest on the host -- exact enough to resolve 1e-7 on fp32 data."""
    a = emb_a.reshape(-1, emb_a.shape[-1]).double()
    b = emb_b.reshape(-1, emb_b.shape[-1]).double()
    Saa, Sbb, Sab = (a.T @ a).cpu().numpy(), (b.T @ b).cpu().numpy(), (a.T @ b).cpu().numpy()
    out = {}
    for d in dims:
        La, Lb = np.linalg.cholesky(Saa[:d, :d]), np.linalg.cholesky(Sbb[:d, :d])
        Cm = np.linalg.solve(La, Sab[:d, :d])
        Cm = np.linalg.solve(Lb, Cm.T).T                       # Qa^T Qb with Q = V L^-T
        cos_min = np.linalg.svd(Cm, compute_uv=False).min()
        out[int(d)] = float(np.sqrt(max(0.0, 1.0 - min(1.0, cos_min) ** 2)))
    return out


def layer_apply(layer: DeviceLayer, x: torch.Tensor, transpose=False, algo=N.ALGO_TCGEN05, k_chunk=0, ws=None):
    """One streaming skinny product (K3/K4): A x or A^T x, x (n, l) fp32."""
    lib = N.load()
    n, l = x.shape
    out = torch.empty_like(x)
    if ws is None:
        ws = torch.empty(lib.n2v2r_apply_workspace_bytes(n, l, k_chunk), dtype=torch.uint8, device=x.device)
    st = layer.c_struct()
    N.check(lib.n2v2r_layer_apply(N.stream_ptr(), C.byref(st), n, l, int(transpose), N.ptr(x), N.ptr(out), algo, k_chunk,
                                  N.ptr(ws), ws.numel()))
    return out


def column_specs(embed_dimensions, distance_metrics):
    """Column order of model.py:73-90: dims outer, metrics inner, cosine skipped at d == 1."""
    names, dims, mets = [], [], []
    for d in embed_dimensions:
        for m in distance_metrics:
            if m == "cosine" and d == 1:
                continue
            if m not in N.METRICS:
                raise NotImplementedError("Unsupported metric")        # model_utils.py:65
            names.append(f"dim-{d}_distance-{m}")
            dims.append(int(d))
            mets.append(N.METRICS[m])
    return names, np.asarray(dims, dtype=np.int32), np.asarray(mets, dtype=np.int32)


def distances(emb: torch.Tensor, col_dims, col_metrics, strategy):
    """K9: scores (n_comparisons, n_cols, n) float64 on the device."""
    lib = N.load()
    T, n, k = emb.shape
    n_comp = T if strategy == N.ONE_VS_REST else T - 1
    n_cols = len(col_dims)
    scores = torch.empty((max(n_comp, 0), n_cols, n), dtype=torch.float64, device=emb.device)
    if n_comp > 0 and n_cols > 0:
        emb = emb.contiguous()
        N.check(lib.n2v2r_distances(N.stream_ptr(), N.ptr(emb), _dtype_code(emb), T, n, k,
                                    col_dims.ctypes.data_as(C.POINTER(C.c_int32)),
                                    col_metrics.ctypes.data_as(C.POINTER(C.c_int32)), n_cols, strategy, N.ptr(scores)))
    return scores


def rank_borda(scores: torch.Tensor, want_positions=False):
    """K10+K11.  scores (n_cols, n) float64 -> (borda (n,) int64, positions (n_cols, n) int32 | None);
    scores (G, n_cols, n) -> (borda (G, n), positions (G, n_cols, n) | None): all G comparisons in one launch."""
    lib = N.load()
    grouped = scores.dim() == 3
    G = scores.shape[0] if grouped else 1
    n_cols, n = scores.shape[-2], scores.shape[-1]
    scores = scores.contiguous()
    ws = torch.empty(lib.n2v2r_rank_workspace_bytes(n, G * n_cols), dtype=torch.uint8, device=scores.device)
    borda = torch.empty((G, n), dtype=torch.int64, device=scores.device)
    pos = torch.empty((G, n_cols, n), dtype=torch.int32, device=scores.device) if want_positions else None
    if G * n_cols > 0:
        N.check(lib.n2v2r_rank_borda(N.stream_ptr(), N.ptr(scores), n, G * n_cols, G, N.ptr(pos), N.ptr(borda),
                                     N.ptr(ws), ws.numel()))
    if grouped:
        return borda, pos
    return borda[0], (pos[0] if pos is not None else None)


def signed_scores(scores: torch.Tensor, prior: torch.Tensor):
    """K11b: scores (cols, n) float64 or int64, prior (n,) float64 -> same shape/dtype, sign taken from the prior
    (kept where prior > 0, negated otherwise; model_utils.py:7-19)."""
    lib = N.load()
    if scores.dtype not in (torch.float64, torch.int64):
        raise TypeError("signed_scores expects float64 or int64 scores")
    scores = scores.contiguous()
    prior = prior.to(torch.float64).contiguous()
    cols, n = scores.shape
    if prior.numel() != n:
        raise ValueError("prior and scores disagree on the number of nodes")
    out = torch.empty_like(scores)
    N.check(lib.n2v2r_signed_scores(N.stream_ptr(), N.ptr(scores), int(scores.dtype == torch.int64), n, cols, N.ptr(prior), N.ptr(out)))
    return out


MAX_EMBED_DIM = 72        # block width l <= 80 (TMEM holds two 3l-column accumulators) minus at least 8 guard vectors


ROWS_GUARD_VECTORS = 8    # guard vectors of the solver block on row-list layers (24 elsewhere)


def default_block_width(k, n, storage=None):
    """Solver block width l: k plus a buffer of guard vectors, a multiple of 16 in [16, 80], < n when possible.
    The dense / tile / e4m3 products cost the same for any l <= 64 (they are bound by the layer stream or the A-operand
    feed), so their block carries 24 guard vectors and needs fewer operator applications; the row-list product is linear
    in l (it is bound by the gather of l x 4 bytes per entry), where 8 guard vectors win: C3 takes 13 applications at
    l = 32 against 11 at l = 48, 102 ms against 120 ms, to the same subspace (scripts/probe_c3_solver.py)."""
    guard = ROWS_GUARD_VECTORS if storage == "rows" else 24
    l = ((k + guard + 15) // 16) * 16
    l = max(16, min(80, l))
    while l > 16 and l >= n:
        l -= 16
    return l


def column_sums(layer: DeviceLayer):
    """Column sums (float64, device) of a DeviceLayer: sum over planes of the bf16 values (model.py:293-294)."""
    if layer.storage != "planes":
        return layer.to_dense().sum(dim=0, dtype=torch.float64)
    return layer.planes[:, :, :layer.n].to(torch.float64).sum(dim=(0, 1))
