"""CLI surface: every reference flag parses, SageMaker quirks are tolerated."""
import pytest

from deepfm_b200.utils.flags import FLAG_SPECS, parse_flags

BASE = ["--field_size", "39", "--feature_size", "117581"]

HVD_FLAGS = ["feature_size", "field_size", "embedding_size", "num_epochs", "batch_size", "log_steps", "learning_rate",
             "l2_reg", "loss_type", "optimizer", "deep_layers", "dropout", "batch_norm", "batch_norm_decay",
             "training_data_dir", "val_data_dir", "checkpoint_dir", "servable_model_dir", "task_type",
             "clear_existing_model", "hosts", "current_host", "pipe_mode", "worker_per_host", "training_channel_name",
             "evaluation_channel_name", "enable_s3_shard", "enable_data_multi_path"]
PS_EXTRA = ["dist_mode", "ps_hosts", "worker_hosts", "job_name", "task_index", "num_threads", "model_dir", "num_GPUs",
            "num_CPUs"]


def test_every_reference_flag_exists_with_reference_defaults():
    for name in HVD_FLAGS + PS_EXTRA:
        assert name in FLAG_SPECS, name
    f, _ = parse_flags(BASE)
    assert (f.embedding_size, f.num_epochs, f.batch_size, f.log_steps) == (32, 10, 64, 1000)
    assert (f.learning_rate, f.l2_reg, f.loss_type, f.optimizer) == (0.0005, 0.0001, "log_loss", "Adam")
    assert f.deep_layer_list == [256, 128, 64] and f.dropout_keep_list == [0.5, 0.5, 0.5]
    assert f.batch_norm is False and f.batch_norm_decay == 0.9 and f.task_type == "train"
    assert f.pipe_mode == 0 and f.worker_per_host == 1 and f.num_threads == 16


def test_environment_is_optional(monkeypatch):
    for k in ("SM_HOSTS", "SM_CURRENT_HOST", "SM_CHANNELS", "SM_NUM_GPUS", "SM_NUM_CPUS"):
        monkeypatch.delenv(k, raising=False)
    f, _ = parse_flags(BASE)
    assert f.hosts == ["localhost"] and f.current_host == "localhost"
    monkeypatch.setenv("SM_HOSTS", '["algo-1", "algo-2"]')
    monkeypatch.setenv("SM_CURRENT_HOST", "algo-2")
    monkeypatch.setenv("SM_NUM_GPUS", "8")
    f, _ = parse_flags(BASE)
    assert f.hosts == ["algo-1", "algo-2"] and f.current_host == "algo-2" and f.num_GPUs == 8


def test_sagemaker_style_booleans_and_unknown_flags():
    f, ignored = parse_flags(BASE + ["--batch_norm", "True", "--enable_s3_shard", "False", "--perform_shuffle", "0",
                                     "--clear_existing_model", "--noenable_data_multi_path", "--pipe_mode=1",
                                     "--optimizer", "ftrl", "--hosts", '["a","b"]'])
    assert f.batch_norm is True and f.enable_s3_shard is False and f.clear_existing_model is True
    assert f.enable_data_multi_path is False and f.pipe_mode == 1 and f.hosts == ["a", "b"]
    assert ignored == ["--perform_shuffle", "0"]


@pytest.mark.parametrize("opt", ["Adam", "Adagrad", "Momentum", "ftrl", "GD", "adam", "FTRL"])
def test_all_optimizers_accepted(opt):
    f, _ = parse_flags(BASE + ["--optimizer", opt])
    assert f.optimizer == opt


def test_bad_values_are_rejected():
    with pytest.raises(SystemExit):
        parse_flags(BASE + ["--task_type", "serve"])
    with pytest.raises(SystemExit):
        parse_flags(["--field_size", "39"])
    with pytest.raises(ValueError):
        parse_flags(BASE + ["--optimizer", "lamb"])
