"""kafka_read_small: scalar read-back through pinned host memory written by a kernel."""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_util as gu
from kafka_inferenceengine_b200 import _lib

pytestmark = pytest.mark.gpu


def test_read_small_round_trip_and_argument_checks():
    for dtype, n in ((torch.float64, 1), (torch.float64, 17), (torch.int32, 2), (torch.int64, 512)):
        t = (torch.arange(n, device=gu.DEV) * 3 + 1).to(dtype)
        got = _lib.read_small(t)
        assert got.dtype == np.dtype(str(dtype).replace("torch.", "")) and np.array_equal(got, t.cpu().numpy())
    lib = _lib.get_lib()
    t = torch.zeros(4, dtype=torch.float64, device=gu.DEV)
    pinned = torch.zeros(4, dtype=torch.float64).pin_memory()
    pageable = np.zeros(4)
    assert lib.kafka_read_small(_lib.ptr(t), C.c_void_p(pinned.data_ptr()), 12, gu.stream()) == -1
    assert lib.kafka_read_small(_lib.ptr(t), None, 16, gu.stream()) == -1
    assert lib.kafka_read_small(_lib.ptr(t), C.c_void_p(pageable.ctypes.data), 16, gu.stream()) > 0   # not pinned
    torch.cuda.synchronize()
    with pytest.raises(ValueError):
        _lib.read_small(torch.zeros(3, dtype=torch.int32, device=gu.DEV))


def test_read_small_is_ordered_after_earlier_work_on_the_stream():
    a = torch.zeros(1 << 24, dtype=torch.float64, device=gu.DEV)
    for k in range(20):
        a += 1.0
        s = a[:2].clone()
        assert np.array_equal(_lib.read_small(s), np.full(2, k + 1.0))
