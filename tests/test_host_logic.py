"""CPU: host-side logic of the drop-in classes that needs no GPU -- the label table behind the label-fused KDE
(membrane.py:273-274 selects points with np.where(np.in1d(labels, ...))) and the lazy result containers."""
import numpy as np
import pytest


@pytest.mark.parametrize("labels", [
    np.array([0, 1, 2, 1, 0, 2, 2, 1]),                       # 0..L-1, all in use: the labels are their own index
    np.array([5, 2, 2, 9, 5, 5, 0, 9], dtype=np.int32),       # gaps: remapped
    np.array([3, 3, 3, 3]),                                   # one label
    np.array(['POPC', 'CHOL', 'POPC', 'DPPC', 'CHOL']),       # strings go through np.unique
    np.array([-1, 4, 4, -1, 7]),                              # negative integers: np.unique, too
    np.array([70000, 1, 70000, 1, 2]),                        # beyond the counting range
])
def test_label_table_equals_np_unique(labels):
    from memsurfer_b200.membrane import _label_table_of
    values, inverse, counts, index32 = _label_table_of(labels, False)
    want_v, want_i, want_c = np.unique(labels, return_inverse=True, return_counts=True)
    assert np.array_equal(values, want_v)
    assert np.array_equal(np.asarray(inverse).reshape(-1), want_i.reshape(-1))
    assert np.array_equal(counts, want_c)
    assert index32.dtype == np.int32 and np.array_equal(index32, want_i.reshape(-1))
    assert np.array_equal(values[index32], labels)


def test_lazy_results_hand_out_arrays_on_every_read_path():
    """LazyResults is a mapping whose values may be downloads in flight: every read path resolves them."""
    from memsurfer_b200 import _lib

    class Fake(_lib.Pending):
        def __init__(self, v):
            self.v, self.n = v, 0

        def result(self):
            self.n += 1
            return self.v

    d = _lib.LazyResults()
    a, b = Fake(np.arange(3)), Fake(np.arange(4))
    d.adopt('a', a)
    d.adopt('b', b)
    d['c'] = np.arange(5)
    assert 'a' in d and len(d) == 3 and sorted(d) == ['a', 'b', 'c']
    assert isinstance(d.raw('a'), _lib.Pending)                # raw() does not wait
    assert np.array_equal(d['a'], np.arange(3)) and a.n == 1
    assert np.array_equal(d['a'], np.arange(3)) and a.n == 1   # resolved once, then a plain array
    assert all(isinstance(v, np.ndarray) for v in d.values())
    assert all(isinstance(v, np.ndarray) for v in dict(d).values())
    assert all(isinstance(v, np.ndarray) for v in {**d}.values())
    other = {}
    other.update(d)
    assert all(isinstance(v, np.ndarray) for v in other.values())
    assert np.array_equal(d.pop('b'), np.arange(4)) and 'b' not in d
    del d['c']
    assert list(d) == ['a']
