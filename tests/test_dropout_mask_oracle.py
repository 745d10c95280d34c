"""CPU: the numpy restatement of the classifier's dropout mask (oracle/dropout_mask.py) is pinned against the
Random123 known-answer vectors for Philox4x32-10, and restates tf.nn.dropout's value rule."""
import numpy as np

from oracle import dropout_mask as dm


def _hex(v):
    return [f"{int(x):08x}" for x in v]


def test_philox4x32_10_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    assert _hex(dm.philox4x32_10([[0, 0, 0, 0]], (0, 0))[0]) == ["6627e8d5", "e169c58d", "bc57ac4c", "9b00dbd8"]
    assert _hex(dm.philox4x32_10([[0xFFFFFFFF] * 4], (0xFFFFFFFF, 0xFFFFFFFF))[0]) == ["408f276d", "41c83b0e", "a20bc7c6", "6d5451fd"]
    assert _hex(dm.philox4x32_10([[0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344]], (0xA4093822, 0x299F31D0))[0]) == [
        "d16cfe09", "94fdcceb", "5001e420", "24126ea1"]


def test_keep_fraction_and_seed_dependence():
    m1 = dm.keep_mask(500, 643, 0.7, 1234567890123)
    m2 = dm.keep_mask(500, 643, 0.7, 1234567890124)
    n = m1.size
    assert abs(m1.mean() - 0.7) < 5 * np.sqrt(0.21 / n)
    assert (m1 != m2).mean() > 0.3  # independent masks differ on 2*0.7*0.3 = 42 % of the elements
    assert (dm.keep_mask(500, 643, 0.7, 1234567890123) == m1).all()
    assert dm.keep_mask(7, 5, 1.0, 3).all()
    # columns and rows are not correlated with the 128-element chunking
    assert abs(m1[:, 0].mean() - 0.7) < 0.12 and abs(m1[0].mean() - 0.7) < 0.1


def test_values_follow_tf_dropout():
    rng = np.random.default_rng(0)
    a = rng.standard_normal((37, 9)).astype(np.float32)
    b = rng.standard_normal((37, 4)).astype(np.float32)
    x = np.concatenate([np.maximum(a, 0), b], 1)
    assert (dm.relu_concat_dropout_forward(a, b, 1.0, 5) == x).all()
    y = dm.relu_concat_dropout_forward(a, b, 0.5, 5)
    m = dm.keep_mask(37, 13, 0.5, 5)
    assert (y[m] == x[m] * np.float32(2)).all() and (y[~m] == 0).all()
    g = rng.standard_normal((37, 13)).astype(np.float32)
    da, db = dm.relu_concat_dropout_backward(g, a, 0.5, 5)
    assert (da == np.where(m[:, :9] & (a > 0), g[:, :9] * np.float32(2), 0)).all()
    assert (db == np.where(m[:, 9:], g[:, 9:] * np.float32(2), 0)).all()
