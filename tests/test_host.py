"""Host-side logic and the C-ABI surface, without a GPU: the shared library loads and exports every symbol the
header declares; the native planner reports the TF variable table; data / flags / settings mirror the reference."""
import os
import re
import subprocess

import numpy as np
import pytest

from util import O, ROOT

from unet_distributed_b200 import _lib, data, parallel, preprocess, settings_dist, synthetic
from unet_distributed_b200.init import initial_weights
from unet_distributed_b200.plan import describe_plan
from unet_distributed_b200.test_dist import build_parser


def test_library_exports_every_declared_symbol():
    protos = _lib.parse_header()
    assert len(protos) >= 45
    lib = _lib.lib()                                  # raises if a declared symbol is missing
    assert lib.unetb200_version() >= 100
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\b(unetb200_\w+)\b", out))
    assert set(protos) <= exported, sorted(set(protos) - exported)
    # nothing is exported that the header does not declare
    assert exported <= set(protos), sorted(exported - set(protos))


def test_library_is_sm100a_tcgen05_tma():
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):   # tcgen05.mma, TMA tensor loads, tcgen05.ld
        assert mnemonic in sass, mnemonic


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    from unet_distributed_b200.plan import UNetPlan
    with pytest.raises(_lib.UnetB200Error):
        UNetPlan(1, 16, 16)


@pytest.mark.parametrize("ups,cin,ncl", [(False, 1, 1), (True, 1, 1), (False, 4, 1), (True, 4, 2)])
def test_native_planner_variable_table(ups, cin, ncl):
    d = describe_plan(2, 32, 32, cin, ncl, ups)
    shapes = O.variable_shapes(ups, cin, ncl)
    assert list(d["variables"]) == list(shapes)                       # TF creation order and names
    for k, (off, shp) in d["variables"].items():
        assert tuple(shp) == tuple(shapes[k]) and off % 4 == 0, k
    assert d["num_params"] == sum(int(np.prod(s)) for s in shapes.values())
    offs = [(off, int(np.prod(s))) for off, s in d["variables"].values()]
    for (o0, n0), (o1, _) in zip(offs, offs[1:]):
        assert o0 + n0 <= o1                                         # no overlap
    assert d["num_params_padded"] >= offs[-1][0] + offs[-1][1]
    # buckets: a partition of the flat buffer, in the order backward completes them
    b = d["buckets"]
    assert len(b) == 5 and b[4][0] == 0 and b[0][1] == d["num_params_padded"]
    assert all(b[i + 1][1] == b[i][0] for i in range(4))
    first = {0: "conv8a/kernel" if ups else "transConv8/kernel", 1: "conv6a/kernel" if ups else "transConv6/kernel",
             2: "conv5a/kernel", 3: "conv3a/kernel", 4: "conv1a/kernel"}
    for i, name in first.items():
        assert b[i][0] == d["variables"][name][0], (i, name)
    assert d["arena_bytes"] > 0


def test_planner_rejects_bad_shapes():
    for args in ((1, 100, 100, 1, 1), (1, 128, 128, 9, 1), (1, 128, 128, 1, 5), (0, 128, 128, 1, 1)):
        with pytest.raises(_lib.UnetB200Error):
            describe_plan(*args)


def test_initial_weights_follow_the_named_initialisers():
    d = describe_plan(1, 16, 16)
    w = initial_weights(d["variables"], seed=7)
    assert sum(a.size for a in w.values()) == 7759521
    assert np.abs(w["conv3b/kernel"]).max() <= np.sqrt(6 / (9 * 128))
    assert np.abs(w["transConv8/kernel"]).max() <= np.sqrt(6 / (4 * 64 + 4 * 128))
    assert np.all(w["conv7a/bias"] == 0)
    w2 = initial_weights(d["variables"], seed=7)
    assert all(np.array_equal(w[k], w2[k]) for k in w)


def test_settings_defaults_are_the_reference_values():
    # settings_dist.py:17-24 of the reference
    assert settings_dist.BATCH_SIZE == 1024 and settings_dist.EPOCHS == 10
    assert settings_dist.LEARNINGRATE == 0.0005 and settings_dist.DECAY_STEPS == 100
    assert settings_dist.LR_FRACTION == 0.2 and settings_dist.CONST_LEARNINGRATE is True
    assert settings_dist.USE_UPSAMPLING is False and settings_dist.MODE == 1
    assert settings_dist.IMG_ROWS == 128 and settings_dist.IN_CHANNEL_NO == 1 and settings_dist.OUT_CHANNEL_NO == 1


def test_flags_accept_both_spellings():
    p = build_parser()
    a = p.parse_args([])
    assert a.batch_size == 1024 and a.learning_rate == 0.0005 and a.const_learningrate is True
    assert a.use_upsampling is False and a.is_sync == 1
    a = p.parse_args(["--learningrate", "0.001", "--use_upsampling=True", "--const_learningrate", "False",
                      "--decay_steps", "150", "--lr_fraction", "0.25", "--batch_size", "128", "--epochs", "3"])
    assert a.learning_rate == 0.001 and a.use_upsampling is True and a.const_learningrate is False
    assert a.decay_steps == 150 and a.lr_fraction == 0.25 and a.batch_size == 128 and a.epochs == 3
    assert p.parse_args(["--learning_rate", "0.002"]).learning_rate == 0.002


def test_get_epoch_contract():
    n, B = 37, 8
    imgs = np.arange(n, dtype=np.float32).reshape(n, 1, 1, 1) * np.ones((1, 4, 4, 1), np.float32)
    msks = -imgs
    ep = data.get_epoch(B, imgs, msks, rng=3)
    assert ep.shape == (4, 2, B, 4, 4, 1)                                # remainder dropped: 37 -> 32
    assert np.array_equal(ep[:, 0], -ep[:, 1])                           # pairs stay together
    ids = ep[:, 0, :, 0, 0, 0].ravel()
    assert len(set(ids.tolist())) == 32 and not np.array_equal(ids, np.arange(32))   # shuffled, no repeats
    assert np.array_equal(data.get_epoch(B, imgs, msks, rng=3), ep)      # seedable
    assert np.array_equal(ep[1, 0], ep[1][0])
    # different channel counts (config 4: 4 modalities, 1 mask channel)
    imgs4 = np.repeat(imgs, 4, axis=3)
    ep4 = data.get_epoch(B, imgs4, msks, rng=3)
    assert ep4[2, 0].shape == (B, 4, 4, 4) and ep4[2, 1].shape == (B, 4, 4, 1) and len(ep4) == 4
    assert np.array_equal(ep4[2, 0][..., 0], -ep4[2, 1][..., 0])


def test_update_channels_and_load_data(tmp_path):
    synthetic.write_npy_dataset(str(tmp_path), n_train=6, n_test=3, H=16, W=16, seed=1)
    imgs, msks = preprocess.load_data(str(tmp_path), "_train")
    assert imgs.shape == (6, 16, 16, 4) and msks.shape == (6, 16, 16, 4)
    i1, m1 = preprocess.update_channels(imgs, msks, 1, 1, 1)
    assert i1.dtype == np.float64 and i1.shape == (6, 16, 16, 1)        # np.zeros default dtype, like the reference
    assert np.array_equal(i1[..., 0], imgs[..., 2]) and np.array_equal(m1[..., 0], msks.sum(axis=3))
    assert set(np.unique(m1)) <= {0.0, 1.0}
    i2, m2 = preprocess.update_channels(imgs, msks, 1, 1, 2)
    assert np.array_equal(i2[..., 0], imgs[..., 0]) and np.array_equal(m2[..., 0], msks[..., 3])
    i3, m3 = preprocess.update_channels(imgs, msks, 1, 1, 3)
    assert np.array_equal(i3[..., 0], imgs[..., 1])
    assert np.array_equal(m3[..., 0], msks[..., 0] + msks[..., 2] + msks[..., 3])
    old = settings_dist.OUT_PATH
    try:
        settings_dist.OUT_PATH = str(tmp_path)
        a, b, c, d = data.load_all_data()
        assert a.shape == (6, 16, 16, 1) and b.shape == (6, 16, 16, 1) and c.shape[0] == 3 and d.shape[0] == 3
    finally:
        settings_dist.OUT_PATH = old


def test_shard_range_is_the_reference_split():
    # test_dist.py:390-392
    for B, n in ((1024, 4), (128, 8), (130, 4), (7, 2)):
        for r in range(n):
            s, e = parallel.shard_range(B, n, r)
            assert s == int(B / n) * r and e - s == int(B / n)


def test_synthetic_slices():
    imgs, msks = synthetic.make_slices(16, 64, 64, 1, seed=816)
    assert imgs.shape == (16, 64, 64, 1) and imgs.dtype == np.float32 and set(np.unique(msks)) <= {0.0, 1.0}
    assert 0 < (msks.reshape(16, -1).sum(1) == 0).sum() < 16            # some empty masks, some not
    i2, _ = synthetic.make_slices(16, 64, 64, 1, seed=816)
    assert np.array_equal(imgs, i2)
    assert os.path.isdir(ROOT)


def test_resident_epochs_match_get_epoch():
    """The device-resident epoch cutter draws the same batches as get_epoch for the same seed (data.py:29-54),
    including the dropped remainder, and hands each replica its shard (test_dist.py:390-394)."""
    import torch
    from unet_distributed_b200.data import ResidentEpochs, get_epoch
    from unet_distributed_b200.parallel import shard_range
    rng = np.random.default_rng(0)
    imgs = rng.standard_normal((37, 8, 8, 4)).astype(np.float32)
    msks = (rng.random((37, 8, 8, 1)) > 0.5).astype(np.float32)
    B = 8
    ep = get_epoch(B, imgs, msks, rng=1234)
    res = ResidentEpochs(B, imgs, msks, torch.device("cpu")).shuffle(1234)
    assert len(res) == len(ep) == 4
    for b in range(len(ep)):
        x, y = res.batch(b)
        np.testing.assert_array_equal(x.numpy(), ep[b, 0])
        np.testing.assert_array_equal(y.numpy(), ep[b, 1])
        for rank in range(2):
            s, e = shard_range(B, 2, rank)
            xs, ys = res.batch(b, s, e)
            np.testing.assert_array_equal(xs.numpy(), ep[b, 0][s:e])
            np.testing.assert_array_equal(ys.numpy(), ep[b, 1][s:e])
    res.shuffle(1235)
    assert not np.array_equal(res.batch(0)[0].numpy(), ep[0, 0])


def test_bench_reference_arm_line(tmp_path):
    """bench.py --impl reference (the CPU restatement timed on the host cores) prints ONE JSON line with the contract's
    keys, and its `config` is the GPU arm's config for the same arguments (same_config in the driver's comparison)."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "1", "--batch", "2", "--size", "32"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
              "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and "cpu_model" in d["cpu_baseline"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"] > 0
    sys.path.insert(0, root)
    import argparse
    import bench
    ns = argparse.Namespace(batch=2, size=32, channels=1, decoder="upsampling", const_learningrate=1, no_graph=False,
                            precision="bf16")
    assert d["config"] == bench.workload_config(ns, 1)


def test_tf32_rounding_of_the_oracle_twin():
    """oracle._tf32_round = PTX cvt.rna.tf32.f32: nearest value with a 10-bit mantissa, ties away from zero."""
    import torch
    x = torch.tensor([1.0, 1.0 + 2.0 ** -11, 1.0 + 2.0 ** -11 - 2.0 ** -20, -(1.0 + 3 * 2.0 ** -11), 3.0e-5, 65504.0])
    r = O._tf32_round(x)
    want = torch.tensor([1.0, 1.0 + 2.0 ** -10, 1.0, -(1.0 + 2.0 ** -9)])
    assert torch.equal(r[:4], want) and float(r[5]) == 65504.0
    assert bool(((r.view(torch.int32) & 0x1FFF) == 0).all())
    assert float(((r - x).abs() / x.abs()).max()) <= 2.0 ** -11 + 1e-9


def test_readme_lists_exactly_the_switches_the_sources_read():
    """Every UNETB200_* environment variable the library, the Python layer, bench.py or the tests read is described in
    README.md, and README.md describes none that no longer exists (a switch removed from the kernels has to leave the
    table too)."""
    import glob
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    documented = set(re.findall(r"UNETB200_[A-Z0-9_]+", open(os.path.join(root, "README.md")).read()))
    files = [f for f in glob.glob(os.path.join(root, "unet_distributed_b200", "**", "*"), recursive=True)
             if f.endswith((".cu", ".cuh", ".h", ".py"))]
    files += [os.path.join(root, "bench.py"), os.path.join(root, "__graft_entry__.py")]
    files += [f for f in glob.glob(os.path.join(root, "tests", "*.py")) if not f.endswith("test_host.py")]
    read = set()
    pat = re.compile(r'getenv\("(UNETB200_[A-Z0-9_]+)"\)|environ(?:\.get)?[\(\[]"(UNETB200_[A-Z0-9_]+)"')
    for f in files:
        for a, b in pat.findall(open(f, errors="ignore").read()):
            read.add(a or b)
    assert read, "no switches found: the pattern is broken"
    assert documented - read == set(), sorted(documented - read)
    assert read - documented == set(), sorted(read - documented)
