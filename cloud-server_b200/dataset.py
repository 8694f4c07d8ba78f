"""Device-resident DataSet: the same ``next_batch`` semantics as the reference's DataSet
(apps/construction/util/construct_distribute.py:466-569) -- shuffle at the first call and at every
epoch wrap, wrap-around batches are tail + head of the reshuffled set -- with the pictures kept in
HBM as float32 rows and each batch assembled by one gather kernel.

The shuffles use numpy's global RNG exactly as the reference does (``numpy.random.shuffle``), on
a host-side index array; only the few hundred batch indices cross PCIe per step."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .context import default_context


class DeviceDataSet:
    def __init__(self, images_u8, labels_onehot, ctx=None):
        """images_u8: uint8 [N, 28, 28(, 1)] (numpy or CUDA tensor); labels: [N, classes] one-hot."""
        self.ctx = ctx or default_context()
        dev = self.ctx.tdev
        if not torch.is_tensor(images_u8):
            images_u8 = torch.from_numpy(np.ascontiguousarray(images_u8))
        images_u8 = images_u8.to(dev).reshape(images_u8.shape[0], -1).contiguous()
        self._num_examples = images_u8.shape[0]
        self._images = self.ctx.normalise(images_u8)                   # astype(float32) * (1/255), :503-504
        lab = labels_onehot if torch.is_tensor(labels_onehot) else torch.from_numpy(np.asarray(labels_onehot, np.float32))
        self._labels = lab.to(dev).float().contiguous()
        assert self._labels.shape[0] == self._num_examples
        self._order = np.arange(self._num_examples)                    # current arrangement (the reference permutes in place)
        self._epochs_completed = 0
        self._index_in_epoch = 0
        self._check_next = True

    num_examples = property(lambda self: self._num_examples)
    epochs_completed = property(lambda self: self._epochs_completed)

    def _gather(self, idx):
        n = len(idx)
        d_idx = torch.from_numpy(np.ascontiguousarray(idx, np.int64)).to(self.ctx.tdev, non_blocking=True)
        L = _lib.lib()
        outs = []
        for src in (self._images, self._labels):
            w = src.shape[1]
            out = torch.empty((n, w), dtype=torch.float32, device=self.ctx.tdev)
            _lib.check(self.ctx.handle, L.nc_gather_rows(self.ctx.handle, C.c_void_p(src.data_ptr()), self._num_examples, w,
                                                         C.c_void_p(d_idx.data_ptr()), n, C.c_void_p(out.data_ptr()),
                                                         self.ctx._stream()))
            outs.append(out)
        if self._check_next:
            # first batch and every epoch wrap (a new permutation): read the kernel's out-of-range flag back (synchronises)
            self._check_next = False
            _lib.check(self.ctx.handle, L.nc_gather_rows_status(self.ctx.handle, self.ctx._stream()))
        return outs[0], outs[1]

    def next_batch(self, batch_size, fake_data=False, shuffle=True):
        start = self._index_in_epoch
        if self._epochs_completed == 0 and start == 0 and shuffle:
            perm0 = np.arange(self._num_examples)
            np.random.shuffle(perm0)
            self._order = self._order[perm0]
        if start + batch_size > self._num_examples:
            self._epochs_completed += 1
            self._check_next = True
            rest = self._order[start:self._num_examples]
            if shuffle:
                perm = np.arange(self._num_examples)
                np.random.shuffle(perm)
                self._order = self._order[perm]
            self._index_in_epoch = batch_size - (self._num_examples - start)
            return self._gather(np.concatenate((rest, self._order[0:self._index_in_epoch])))
        self._index_in_epoch += batch_size
        return self._gather(self._order[start:self._index_in_epoch])
