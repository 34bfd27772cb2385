"""Row-band sharding of one frame over devices - the host-side statement of the library's RowMap
(csrc/renderer.h) that callers use to know which rows a rank owns.

The reference deals 20-row work items of a column to 8 goroutines (raytracer.go:653-672); here the
image is cut into bands of BAND_ROWS rows dealt round-robin to the parts, and a part renders its
bands as one contiguous run of *virtual* rows.  Pixels are independent, so there is no exchange
step: the gather is either stores into one frame (CUDA IPC / peer access) or a SUM of frames that
are zero outside their own bands."""
import numpy as np

BAND_ROWS = 8


def actual_row(v, part, n_parts, band_rows=BAND_ROWS):
    """Virtual row v of `part` -> image row (RowMap::actual)."""
    if n_parts == 1:
        return v
    return ((v // band_rows) * n_parts + part) * band_rows + v % band_rows


def n_rows(part, n_parts, height, band_rows=BAND_ROWS):
    """Number of image rows `part` owns (RowMap::count)."""
    if n_parts == 1:
        return height
    full, rest = divmod(height, band_rows)
    rows = (full // n_parts + (1 if full % n_parts > part else 0)) * band_rows
    if rest and full % n_parts == part:
        rows += rest
    return rows


def rows_of(part, n_parts, height, band_rows=BAND_ROWS):
    """Image rows owned by `part`, in the order the device renders them."""
    return np.array([actual_row(v, part, n_parts, band_rows) for v in range(n_rows(part, n_parts, height, band_rows))], dtype=np.int64)


def own_row_blocks(part, n_parts, height, band_rows=BAND_ROWS):
    """The rows of `part` as gmlrt_download_rows copies them: one strided 2-D copy of the full bands plus the partial band
    at the bottom when `part` owns it -> ((first_row, pitch_rows, rows_per_band, n_bands), (tail_first_row, tail_rows))."""
    full, rest = divmod(height, band_rows)
    mine = full // n_parts + (1 if full % n_parts > part else 0)
    tail = (full * band_rows, rest) if rest and full % n_parts == part else (0, 0)
    return (part * band_rows, n_parts * band_rows, band_rows, mine), tail


def copy_own_rows(dst, src, part, n_parts, band_rows=BAND_ROWS):
    """Host statement of gmlrt_download_rows: rows of `part` from `src` into `dst` ([H, ...] arrays), others untouched."""
    (first, pitch, br, nb), (t0, tn) = own_row_blocks(part, n_parts, dst.shape[0], band_rows)
    for b in range(nb):
        y = first + b * pitch
        dst[y:y + br] = src[y:y + br]
    dst[t0:t0 + tn] = src[t0:t0 + tn]
    return dst


def owner_of(y, n_parts, band_rows=BAND_ROWS):
    return (y // band_rows) % n_parts


def gather_sum(frame, dist, dst=0):
    """Gather frames that are zero outside their own bands onto `dst` with one reduce(SUM) (torch.distributed,
    any backend).  uint8 SUM of disjoint bands cannot overflow."""
    dist.reduce(frame, dst=dst, op=dist.ReduceOp.SUM)
    return frame
