"""Reference arm of bench.py: runs the UNMODIFIED reference (copied under baseline/_ref) through its own public
C++ API (Sequential / SequentialBuilder model zoo, Adam, LogSoftmaxCrossEntropy, the train_class_epoch loop of
include/nn/train.hpp:108-170) on synthetic data.  See baseline/README.md for how the subset build is made."""
import json
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "_ref", "bin", "ref_bench")


def run(args):
    if getattr(args, "device", "cuda") == "cpu":
        # BASELINE config 1: the reference's MNIST CNN on the CPU (serial parallel_for in the offline subset build)
        cpu_bin = os.path.join(HERE, "_ref", "bin", "ref_mnist_cpu")
        if not os.path.exists(cpu_bin) or getattr(args, "model", "") != "mnist_cnn":
            print(json.dumps({"impl": "reference", "unavailable": "CPU reference arm exists for --model mnist_cnn only "
                              "(baseline/build_ref.sh builds it)"}))
            return 0
        r = subprocess.run([cpu_bin, str(args.batch), str(args.steps), str(args.warmup)], capture_output=True, text=True)
        line = [l for l in r.stdout.splitlines() if l.startswith("{")]
        print(line[-1] if r.returncode == 0 and line else
              json.dumps({"impl": "reference", "unavailable": "reference CPU harness failed"}))
        return 0
    if not os.path.exists(BIN):
        print(json.dumps({"impl": "reference", "unavailable": "reference harness not built (baseline/build_ref.sh)"}))
        return 0
    if args.gpus != 1:
        print(json.dumps({"impl": "reference", "unavailable":
                          "the reference has no data parallelism and its pipeline needs asio+TBB, absent offline; "
                          "only the single-GPU reference arm exists"}))
        return 0
    cmd = [BIN, str(args.batch), str(args.steps), str(args.warmup)]
    model = getattr(args, "model", "resnet50_tiny_imagenet")
    if model != "resnet50_tiny_imagenet":      # other BASELINE.json configs: same harness, another zoo model
        cmd.append(model)
    # The reference's DeviceManager::getDevice(DeviceType::GPU) returns the first match of an unordered_map
    # (/root/reference/src/device/device_manager.cpp:83-90): with several GPUs visible it may pick one its context was
    # not made for.  Show it exactly one device (environment only -- the reference itself is untouched).
    env = dict(os.environ)
    vis = env.get("CUDA_VISIBLE_DEVICES", "")
    env["CUDA_VISIBLE_DEVICES"] = (vis.split(",")[0] if vis else "0")
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    line = None
    for l in r.stdout.splitlines():
        if l.startswith("{"):
            line = l
    if r.returncode != 0 or line is None:
        print(json.dumps({"impl": "reference", "unavailable": "reference harness failed: " +
                          (r.stderr.strip().splitlines() or ["no output"])[-1][:200]}))
        return 0
    print(same_keys(line))
    return 0


L2_NOTE = ("per-step working set (activations, >10 GB at batch 256) is far larger than the 126 MB L2; "
           "inputs rotate over 4 distinct batches")


def same_keys(line):
    """Both arms print the same `config` key set (bench.py's), with the values of a single-device run."""
    try:
        d = json.loads(line)
    except ValueError:
        return line
    c = d.get("config", {})
    full = {"model": c.get("model"), "global_batch": c.get("global_batch"), "per_gpu_batch": c.get("per_gpu_batch"),
            "seq_len": None, "image": c.get("image"), "classes": c.get("classes"), "parallelism": "single",
            "optimizer": c.get("optimizer"), "loss": c.get("loss"), "microbatches": 1, "pipeline_stages": 1,
            "dp_replicas": 1, "l2": L2_NOTE}
    d["config"] = full
    return json.dumps(d)
