"""GPU: `:wf2model` checkpoints of device tensors, mirroring the reference's own
save-and-restore test (`/root/reference/source/vm/t/utils.lisp:20-45`): save a model, build a fresh one,
load, the forward result is identical."""
import importlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_save_and_restore_model(gpu, tmp_path):
    nn = importlib.import_module("cl-waffe2_b200.nn")
    sd = importlib.import_module("cl-waffe2_b200.state_dict")
    dims = [20, 10, 5]
    X = np.arange(20, dtype=np.float32).reshape(1, 20)
    a = nn.MLPSequence(dims, "relu", seed=1)
    want = a(gpu.from_numpy(X)).to_numpy()
    path = str(tmp_path / "model_weight_restore_tester.wf2model")
    assert sd.save_weights(path, sd.make_state_dict(a.named_parameters()))
    # layout of the reference's writer: parameters.json maps each key to "<dir>/<key>.npy", a 1-D float32 array
    table = json.load(open(os.path.join(path, "parameters.json")))
    assert sorted(table) == ["param:linearlayer.0.bias", "param:linearlayer.0.weights",
                             "param:linearlayer.1.bias", "param:linearlayer.1.weights"]
    for key, file in table.items():
        assert file == f"{path}/{key}.npy"
        arr = np.load(file)
        assert arr.dtype == np.float32 and arr.ndim == 1
    np.testing.assert_array_equal(np.load(table["param:linearlayer.0.weights"]), a.parameters()[0].to_numpy().ravel())
    b = nn.MLPSequence(dims, "relu", seed=2)                      # different weights
    assert not np.array_equal(b(gpu.from_numpy(X)).to_numpy(), want)
    failed = sd.load_from_state_dict(b.named_parameters(), sd.load_weights(path))
    assert failed == []
    np.testing.assert_array_equal(b(gpu.from_numpy(X)).to_numpy(), want)


def test_optimizer_state_round_trip_resumes_training(gpu, tmp_path):
    # save after 3 Adam steps with optimizers=T, restore into a fresh model + optimizer, continue: the
    # next losses equal those of the uninterrupted run bit for bit
    nn = importlib.import_module("cl-waffe2_b200.nn")
    sd = importlib.import_module("cl-waffe2_b200.state_dict")
    rng = np.random.default_rng(0)
    dims = [12, 8, 4]
    X = rng.uniform(0, 1, (16, 12)).astype(np.float32)
    Y = np.eye(4, dtype=np.float32)[rng.integers(4, size=16)]

    def make(seed):
        m = nn.MLPSequence(dims, "relu", seed=seed)
        opt = nn.Adam(m.arena, lr=1e-2)
        st = nn.CompiledStep(m, "softmax-cross-entropy-sum", opt, X.shape, Y.shape, use_graph=False)
        st.set_inputs(X, Y)
        return m, opt, st

    m1, o1, s1 = make(5)
    for _ in range(3):
        s1.step()
    path = str(tmp_path / "ckpt.wf2model")
    sd.save_weights(path, sd.make_state_dict(m1.named_parameters(), o1, optimizers=True))
    table = json.load(open(os.path.join(path, "parameters.json")))
    assert table["optimizer:linearlayer.0.weights.adam.single-float.lr"] == pytest.approx(1e-2)
    assert table["optimizer:linearlayer.1.bias.adam.(integer 0 4611686018427387903).n"] == 3
    assert table["optimizer:linearlayer.0.bias.adam.cudatensor.m"].endswith("optimizer:linearlayer.0.bias.adam.cudatensor.m.npy")
    cont = []
    for _ in range(3):
        s1.step()
        cont.append(s1.loss_value())
    m2, o2, s2 = make(6)
    failed = sd.load_from_state_dict(m2.named_parameters(), sd.load_weights(path), o2)
    assert failed == [] and o2.n == 3
    resumed = []
    for _ in range(3):
        s2.step()
        resumed.append(s2.loss_value())
    assert resumed == cont


def test_load_reports_unknown_keys_and_rejects_size_mismatch(gpu, tmp_path):
    nn = importlib.import_module("cl-waffe2_b200.nn")
    sd = importlib.import_module("cl-waffe2_b200.state_dict")
    a = nn.MLPSequence([6, 3], "relu", seed=1)
    d = sd.make_state_dict(a.named_parameters())
    d["param:conv2d.0.weight"] = gpu.make_tensor([2, 2])
    path = str(tmp_path / "x.wf2model")
    sd.save_weights(path, d)
    assert sd.load_from_state_dict(a.named_parameters(), sd.load_weights(path)) == ["param:conv2d.0.weight"]
    b = nn.MLPSequence([6, 4], "relu", seed=1)
    with pytest.raises(ValueError):
        sd.load_from_state_dict(b.named_parameters(), sd.load_weights(path))
