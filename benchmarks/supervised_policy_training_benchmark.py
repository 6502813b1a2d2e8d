"""The reference's benchmarks/supervised_policy_training_benchmark.py:1-20 on this package: a 12-layer, 128-filter policy
on ['board', 'ones', 'turns_since'], run_training for 5 epochs at minibatch 32 and learning rate 0.01, under cProfile.
(The reference script cannot run as written: it passes six positionals to run_training(cmd_line_args=None) and names a
dataset that is not in its tree.)  The dataset here is converted on the spot from the game records under tests/golden/sgf.
Needs a B200.

    python benchmarks/supervised_policy_training_benchmark.py [out_dir]   ->  supervised_policy_training_bench_results.prof
"""
import cProfile
import glob
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alphago_b200.game_converter import game_converter
from alphago_b200.models import CNNPolicy
from alphago_b200.supervised_policy_trainer import run_training

features = ["board", "ones", "turns_since"]
work = sys.argv[1] if len(sys.argv) > 1 else tempfile.mkdtemp()
os.makedirs(work, exist_ok=True)
model_file, data_file, out_dir = (os.path.join(work, n) for n in ("model.json", "debug_feature_planes.hdf5", "training_results"))
CNNPolicy(features, filters_per_layer=128, layers=12).save_model(model_file)
game_converter(features).sgfs_to_hdf5(sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "sgf", "*.sgf"))), data_file, bd_size=19)

profiler = cProfile.Profile()
t0 = time.time()
meta = profiler.runcall(run_training, [model_file, data_file, out_dir, "--epochs", "5", "--minibatch", "32", "--learning-rate", "0.01"])
seconds = time.time() - t0
profiler.dump_stats("supervised_policy_training_bench_results.prof")
print(json.dumps({"epochs": len(meta["epochs"]), "seconds": round(seconds, 2), "last_epoch": meta["epochs"][-1],
                  "profile": "supervised_policy_training_bench_results.prof", "out_directory": out_dir}))
