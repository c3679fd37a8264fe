"""Device selection and the scalar (one row at a time) bridge to libacx.so."""
import os

import numpy as np
import torch

from acx import _lib
from acx.batch import _prep

MODE_CYCLICAL, MODE_NO_MOVE, MODE_RAW_MOVE = 1, 2, 4


def device():
    if not torch.cuda.is_available():
        raise _lib.AcxError("ac_solver (acx build) needs a CUDA device: the AC-move engine has no CPU path")
    return torch.device(os.environ.get("ACX_DEVICE", "cuda:%d" % torch.cuda.current_device()))


def rows_primitive(row, m, move, mode):
    """One row through acx_rows_primitive.  Returns (row int8, [len0, len1], status)."""
    dev = device()
    stream = _prep(dev)
    a = np.ascontiguousarray(row, dtype=np.int8).reshape(1, 2 * m)
    t = torch.from_numpy(a).to(dev)
    out = torch.empty_like(t)
    mv = torch.tensor([int(move)], dtype=torch.uint8, device=dev)
    lens = torch.empty((1, 2), dtype=torch.uint8, device=dev)
    st = torch.empty((1,), dtype=torch.uint8, device=dev)
    rc = _lib.lib().acx_rows_primitive(t.data_ptr(), mv.data_ptr(), out.data_ptr(), lens.data_ptr(), st.data_ptr(),
                                       1, m, mode, stream)
    _lib.check(rc, "acx_rows_primitive")
    packed = torch.cat([out.view(torch.uint8).flatten(), lens.flatten(), st]).cpu().numpy()
    row_out = packed[: 2 * m].view(np.int8).copy()
    return row_out, [int(packed[2 * m]), int(packed[2 * m + 1])], int(packed[2 * m + 2])
