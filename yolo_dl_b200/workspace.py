"""Per-module scratch of the C ABI (yh_workspace_create / destroy)."""
import ctypes

import torch

from . import _lib


class Workspace:
    """Grows on demand; one per module instance and device (SURVEY.md §8b threading: no global
    mutable state, a module used from several threads needs one workspace per thread)."""

    def __init__(self):
        self._lib = _lib.load()
        self.handle = None
        self.key = None
        self._rows = 0

    def ensure(self, device: torch.device, batch: int, instances: int, cand_per_image: int = 1):
        if not (isinstance(device, torch.device) and device.type == "cuda"):
            raise _lib.YoloHotError(_lib.YH_ERR_CUDA, "yolohot needs a CUDA device; there is no CPU fallback")
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self.key is not None:
            d, b, n, c = self.key
            if d == idx and b >= batch and n >= instances and c >= cand_per_image:
                return self.handle
            batch, instances, cand_per_image = max(b, batch), max(n, instances), max(c, cand_per_image)
            self.close()
        h = ctypes.c_void_p()
        _lib.check(self._lib.yh_workspace_create(idx, batch, instances, cand_per_image, ctypes.byref(h)))
        self.handle = h
        self.key = (idx, batch, instances, cand_per_image)
        self._rows = 0
        return h

    def reserve_rows(self, capacity: int, num_classes: int):
        """yh_workspace_reserve_rows: staging of the matched-row gradients for `capacity` rows.
        The loss entry points never allocate (they fail with YH_ERR_CAPACITY instead); this is the
        caller-side reserve, grown geometrically so that a training loop reallocates (and
        synchronises) only a handful of times."""
        need = capacity * (5 + num_classes)
        if need <= self._rows:
            return
        cap = max(capacity, 2 * (self._rows // (5 + num_classes)))
        _lib.check(self._lib.yh_workspace_reserve_rows(self.handle, cap, num_classes))
        self._rows = cap * (5 + num_classes)

    def read_counts(self, stream, M_dev=None, K_dev=None):
        """The one synchronising read (yh_read_counts); raises on a buffer overflow."""
        m, k = ctypes.c_int32(0), ctypes.c_int32(0)
        st = self._lib.yh_read_counts(self.handle, M_dev.data_ptr() if M_dev is not None else None,
                                      ctypes.byref(m), K_dev.data_ptr() if K_dev is not None else None,
                                      ctypes.byref(k), stream)
        _lib.check(st)
        return m.value, k.value

    def close(self):
        if self.handle is not None:
            self._lib.yh_workspace_destroy(self.handle)
            self.handle = None
            self.key = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
