"""The CIFAR oracle restatement against golden vectors made by the reference's own
classes (oracle/make_golden.py) -- this is what pins the oracle."""
import os

import numpy as np
import torch

from oracle import cifar_oracle as co


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


def test_init_matches_reference_digest(golden_dir):
    g = _load(golden_dir, "cifar_b8.npz")
    sd = co.init_state_dict(int(g["weight_seed"]))
    assert co.state_dict_digest(sd) == str(g["digest"])
    assert {k: tuple(v.shape) for k, v in sd.items()} == co.PARAM_SHAPES


def test_parts_bit_exact_vs_reference(golden_dir):
    torch.set_num_threads(1)                       # golden made single-threaded
    g = _load(golden_dir, "cifar_b8.npz")
    sd = co.init_state_dict(int(g["weight_seed"]))
    x = co.make_input(8, int(g["input_seed"]))
    assert np.array_equal(x.numpy(), g["x"])
    p0 = co.part0(x, sd)
    assert p0.shape == (8, 4096)
    np.testing.assert_allclose(p0.numpy(), g["part0"], rtol=0, atol=1e-6)
    p1 = co.part1(torch.from_numpy(g["part0"]), sd)
    np.testing.assert_allclose(p1.numpy(), g["probs"], rtol=0, atol=1e-7)
    assert np.array_equal(p1.argmax(1).numpy(), g["probs"].argmax(1))
    np.testing.assert_allclose(co.full(x, sd).numpy(), g["probs"], rtol=0, atol=1e-6)


def test_flatten_order_is_nchw(golden_dir):
    """fc1's weights assume index = c*64 + h*8 + w (cifar_model_parts.py:41)."""
    sd = co.init_state_dict(1234)
    x = co.make_input(2, 3)
    import torch.nn.functional as F
    a = F.max_pool2d(F.relu(F.conv2d(x, sd["conv1.weight"], sd["conv1.bias"], 1, 1)), 2, 2)
    a = F.max_pool2d(F.relu(F.conv2d(a, sd["conv2.weight"], sd["conv2.bias"], 1, 1)), 2, 2)
    flat = co.part0(x, sd)
    assert flat[1, 5 * 64 + 3 * 8 + 2] == a[1, 5, 3, 2]


def test_probabilities_sum_to_one_and_edge_batches():
    sd = co.init_state_dict(1234)
    for b in (1, 3):
        p = co.full(co.make_input(b, 11), sd)
        assert p.shape == (b, 10)
        np.testing.assert_allclose(p.sum(1).numpy(), np.ones(b), atol=1e-6)


def test_recorded_grpc_run_equals_in_process(golden_dir):
    """the two-process localhost run of the unmodified node.py returns the
    in-process result bit for bit, and its status strings have the documented form."""
    torch.set_num_threads(1)
    g = _load(golden_dir, "cifar_grpc.npz")
    sd = co.init_state_dict(int(g["weight_seed"]))
    for b in (1, 4):
        x = torch.from_numpy(g[f"x{b}"])
        np.testing.assert_allclose(co.full(x, sd).numpy(), g[f"probs{b}"], rtol=0, atol=1e-6)
        flat = int(np.argmax(g[f"probs{b}"]))
        assert str(g[f"status{b}"]) == (f"[node1] Forwarded. Next node status: "
                                         f"[node2] Processing complete. Prediction: {flat}")
    assert bool(g["healthy"]) is True
    assert str(g["confirmation"]) == "[node1] got msg 'hi'"
