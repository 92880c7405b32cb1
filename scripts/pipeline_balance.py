"""How well can ResNet-50 be cut into S pipeline stages?  Measures forward+backward GPU time per top-level layer at the
microbatch size the pipeline runs, partitions with BalancedPartitioner (min-max over contiguous ranges) and prints the
stage sums: max / mean is the factor by which the slowest stage stretches the steady state of a 1F1B schedule."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200.models import MODEL_ZOO  # noqa: E402
from dcnn_b200.partitioner import BalancedPartitioner, measure_layer_costs  # noqa: E402

model_name = sys.argv[1] if len(sys.argv) > 1 else "resnet50_tiny_imagenet"
stages = int(sys.argv[2]) if len(sys.argv) > 2 else 4
hw = 32 if "cifar" in model_name else 64
for mb in [int(v) for v in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["64", "128", "256"])]:
    m = MODEL_ZOO[model_name]()
    m.set_seed(1)
    m.set_device("cuda:0")
    m.initialize()
    costs = measure_layer_costs(m, [mb, 3, hw, hw], warmup=2, iters=3)
    parts = BalancedPartitioner(costs=costs).get_partitions(m.get_layers(), stages)
    sums = [sum(costs[p.start_layer:p.end_layer]) for p in parts]
    mean = sum(sums) / len(sums)
    print(f"{model_name} microbatch {mb}: layers {len(costs)}, total {sum(costs):.3f} ms, stages "
          f"{[(p.start_layer, p.end_layer) for p in parts]} sums {[round(s, 3) for s in sums]} ms, "
          f"max/mean {max(sums) / mean:.3f}, largest single layer {max(costs):.3f} ms "
          f"({max(costs) / sum(costs) * 100:.1f} % of the total)")
    del m
    torch.cuda.empty_cache()
