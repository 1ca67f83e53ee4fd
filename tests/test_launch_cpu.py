"""The local stand-ins for the SageMaker launchers and set_dist_env (C9)."""
import json
import os
import subprocess
import sys
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from deepfm_b200 import launch  # noqa: E402
from deepfm_b200.data import tfrecord as T  # noqa: E402
from deepfm_b200.parallel.env import set_dist_env  # noqa: E402


def test_set_dist_env_modes(monkeypatch):
    monkeypatch.delenv("TF_CONFIG", raising=False)
    assert set_dist_env(0, "p:1", "w:1", "worker", 0) is None and "TF_CONFIG" not in os.environ
    c = set_dist_env(1, "p0:2223,p1:2223", "w0:2222,w1:2222,w2:2222", "worker", 0)
    assert c == {"cluster": {"chief": ["w0:2222"], "ps": ["p0:2223", "p1:2223"], "worker": ["w1:2222", "w2:2222"]},
                 "task": {"type": "chief", "index": 0}}
    assert json.loads(os.environ["TF_CONFIG"]) == c
    assert set_dist_env(1, "p0:1", "w0:1,w1:1", "worker", 1)["task"] == {"type": "worker", "index": 0}
    assert set_dist_env(1, "p0:1", "w0:1,w1:1", "ps", 0)["task"] == {"type": "ps", "index": 0}
    c = set_dist_env(2, "p0:1", "w0:1,w1:1,w2:1,w3:1", "worker", 3)
    assert c["cluster"] == {"chief": ["w0:1"], "ps": ["p0:1"], "worker": ["w2:1", "w3:1"]}      # no worker is dropped
    assert c["task"] == {"type": "worker", "index": 1}
    assert set_dist_env(2, "p0:1", "w0:1,w1:1,w2:1", "worker", 1)["task"] == {"type": "evaluator", "index": 0}


def _args(**kw):
    d = dict(distribution="mpi", instance_count=1, processes_per_host=1, input_mode="File", hyperparameters=[],
             channels=[], pipe_root="", master_port=29555, dry_run=True)
    d.update(kw)
    return types.SimpleNamespace(**d)


def test_mpi_launch_plan():
    procs = launch.build_commands(_args(processes_per_host=4, hyperparameters=["batch_size=1024", "field_size=39"],
                                        channels=["training=/d/tr", "evaluation=/d/va"]))
    assert len(procs) == 4
    for r, p in enumerate(procs):
        e, a = p["env"], p["argv"]
        assert (e["RANK"], e["LOCAL_RANK"], e["WORLD_SIZE"]) == (str(r), str(r), "4")
        assert json.loads(e["SM_CHANNELS"]) == ["evaluation", "training"] and e["SM_CHANNEL_TRAINING"] == "/d/tr"
        assert a[a.index("--batch_size") + 1] == "1024" and a[a.index("--training_data_dir") + 1] == "/d/tr"
        assert a[a.index("--worker_per_host") + 1] == "4" and a[a.index("--dist") + 1] == "hvd"


def test_parameter_server_launch_plan():
    procs = launch.build_commands(_args(distribution="parameter_server", instance_count=2,
                                        channels=["training=/d/tr", "evaluation=/d/va"]))
    roles = [(json.loads(p["env"]["TF_CONFIG"])["task"]["type"], json.loads(p["env"]["TF_CONFIG"])["task"]["index"])
             for p in procs]
    assert roles == [("ps", 0), ("ps", 1), ("chief", 0), ("worker", 0)]
    cl = json.loads(procs[0]["env"]["TF_CONFIG"])["cluster"]
    assert len(cl["ps"]) == 2 and len(cl["chief"]) == 1 and len(cl["worker"]) == 1


def _write_split(tmp_path, val_tfrecords):
    recs = list(T.iter_records(val_tfrecords))[:1600]
    (tmp_path / "train").mkdir()
    (tmp_path / "val").mkdir()
    for name, part in (("train/tr_0.tfrecords", recs[:600]), ("train/tr_1.tfrecords", recs[600:1200]),
                       ("val/va_0.tfrecords", recs[1200:])):
        with T.TFRecordWriter(str(tmp_path / name)) as w:
            for r in part:
                w.write(r)


HP = ["field_size=39", "feature_size=117581", "deep_layers=16,8", "dropout=0.9,0.9", "embedding_size=8",
      "batch_size=100", "backend=eager", "log_steps=2", "num_epochs=1"]


def _launch(argv, timeout=420):
    env = dict(os.environ, PYTHONPATH=ROOT, CUDA_VISIBLE_DEVICES="")
    for k in ("SM_HOSTS", "SM_CURRENT_HOST", "SM_CHANNELS", "RANK", "WORLD_SIZE", "TF_CONFIG"):
        env.pop(k, None)
    return subprocess.run([sys.executable, "-m", "deepfm_b200.launch"] + argv, env=env, capture_output=True, text=True,
                          timeout=timeout)


def test_launch_mpi_two_ranks_file_mode(tmp_path, val_tfrecords):
    """NB-HVD file mode in miniature: two ranks, synchronous data parallel over gloo, rank 0 evaluates and saves."""
    _write_split(tmp_path, val_tfrecords)
    r = _launch(["--distribution", "mpi", "--processes_per_host", "2", "--master_port", "29571",
                 "--hyperparameters"] + HP + [f"checkpoint_dir={tmp_path / 'ckpt'}", "enable_s3_shard=False",
                 "--channels", f"training={tmp_path / 'train'}", f"evaluation={tmp_path / 'val'}"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "current rank is 0 of 2" in r.stdout and "'auc'" in r.stdout
    assert os.path.exists(tmp_path / "ckpt" / "checkpoint")


def test_launch_pipe_mode_single_rank(tmp_path, val_tfrecords):
    """NB-HVD pipe mode in miniature: the channels arrive as FIFOs <root>/<channel>_<epoch>."""
    _write_split(tmp_path, val_tfrecords)
    r = _launch(["--distribution", "mpi", "--processes_per_host", "1", "--input_mode", "Pipe", "--master_port", "29572",
                 "--pipe_root", str(tmp_path), "--hyperparameters"] + HP +
                [f"checkpoint_dir={tmp_path / 'ckpt'}", "training_channel_name=training",
                 "--channels", f"training={tmp_path / 'train'}", f"evaluation={tmp_path / 'val'}"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "'auc'" in r.stdout and os.path.exists(tmp_path / "ckpt" / "checkpoint")


def test_launch_pipe_mode_two_ranks_with_uneven_channels(tmp_path, val_tfrecords):
    """NB-HVD pipe mode (256-301): one training channel per local worker (HVD.py:445), and the channels do not hold the
    same number of records.  The reference hangs in Horovod when one rank runs out; here both ranks stop
    on the same step and the job finishes: evaluation, checkpoint, export."""
    recs = list(T.iter_records(val_tfrecords))[:1300]
    for d, part in (("t0", recs[:600]), ("t1", recs[600:950]), ("val", recs[950:])):
        (tmp_path / d).mkdir()
        with T.TFRecordWriter(str(tmp_path / d / ("va_0.tfrecords" if d == "val" else "tr_0.tfrecords"))) as w:
            for r in part:
                w.write(r)
    r = _launch(["--distribution", "mpi", "--processes_per_host", "2", "--input_mode", "Pipe", "--master_port", "29575",
                 "--pipe_root", str(tmp_path / "pipes"), "--hyperparameters"] + HP +
                [f"checkpoint_dir={tmp_path / 'ckpt'}", "log_steps=1", "enable_data_multi_path=True",
                 "--channels", f"training0={tmp_path / 't0'}", f"training1={tmp_path / 't1'}",
                 f"evaluation={tmp_path / 'val'}"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "current rank is 1 of 2" in r.stdout and "'auc'" in r.stdout
    assert "step = 3 " in r.stdout and "step = 4 " not in r.stdout        # 350 records = 3 full batches of 100
    assert os.path.exists(tmp_path / "ckpt" / "checkpoint")


def test_launch_parameter_server_two_instances(tmp_path, val_tfrecords):
    """NB-PS in miniature: two 'instances', each with one PS task and one worker task wired by TF_CONFIG."""
    _write_split(tmp_path, val_tfrecords)
    r = _launch(["--distribution", "parameter_server", "--instance_count", "2", "--master_port", "29573",
                 "--hyperparameters"] + HP + [f"model_dir={tmp_path / 'model'}",
                 "--channels", f"training={tmp_path / 'train'}", f"evaluation={tmp_path / 'val'}"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "[ps 0] serving" in r.stdout and "[ps 1] serving" in r.stdout and "'auc'" in r.stdout
    assert os.path.exists(tmp_path / "model" / "checkpoint")


def test_parameter_server_infer_export_and_pipe(tmp_path, val_tfrecords):
    """PS.py:137-151 (pipe input) and PS.py:522-551 (eval / infer -> pred.txt / export-only) in parameter-server mode:
    train through FIFOs, then infer and export from the checkpoint the chief wrote."""
    _write_split(tmp_path, val_tfrecords)
    recs = list(T.iter_records(val_tfrecords))[1200:1500]
    with T.TFRecordWriter(str(tmp_path / "val" / "te_0.tfrecords")) as w:
        for r in recs:
            w.write(r)
    common = HP + [f"model_dir={tmp_path / 'model'}"]
    r = _launch(["--distribution", "parameter_server", "--instance_count", "1", "--input_mode", "Pipe",
                 "--pipe_root", str(tmp_path), "--master_port", "29574", "--hyperparameters"] + common +
                ["training_channel_name=training", "evaluation_channel_name=evaluation",
                 "--channels", f"training={tmp_path / 'train'}", f"evaluation={tmp_path / 'val'}"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert os.path.exists(tmp_path / "model" / "checkpoint")
    r = _launch(["--distribution", "parameter_server", "--instance_count", "1", "--master_port", "29575",
                 "--hyperparameters"] + common + ["task_type=infer",
                 "--channels", f"training={tmp_path / 'train'}", f"evaluation={tmp_path / 'val'}"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    preds = [float(x) for x in open(tmp_path / "val" / "pred.txt").read().split()]
    assert len(preds) == 300 and all(0.0 <= p <= 1.0 for p in preds)
    r = _launch(["--distribution", "parameter_server", "--instance_count", "1", "--master_port", "29576",
                 "--hyperparameters"] + common + ["task_type=export", f"servable_model_dir={tmp_path / 'servable'}",
                 "--channels", f"training={tmp_path / 'train'}", f"evaluation={tmp_path / 'val'}"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert any((tmp_path / "servable").iterdir())
