"""What the reference's benchmarks/preprocessing_benchmark.py:1-16 measures -- game_converter.convert_game over one
186-move record with 8 features, under cProfile -- on this package, plus the bulk path the converter really uses
(pack_game + one encoder launch).  Needs a B200: the planes come from the CUDA encoder.

    python benchmarks/preprocessing_benchmark.py [game.sgf]      ->  bench_results.prof + one JSON line
"""
import cProfile
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alphago_b200.game_converter import game_converter

FEATURES = ["board", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after", "sensibleness", "zeros"]
record = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "sgf", "Lee-Sedol-vs-AlphaGo-20160309.sgf")
converter = game_converter(FEATURES)
sum(1 for _ in converter.convert_game(record, 19))           # first call: library load, CUDA context


def per_position():
    return sum(1 for _ in converter.convert_game(record, 19))


profiler = cProfile.Profile()
t0 = time.time()
pairs = profiler.runcall(per_position)
t_gen = time.time() - t0
profiler.dump_stats("bench_results.prof")
t0 = time.time()
packed, actions, error = converter.pack_game(record, 19)
planes = converter.encode_packed_positions(packed, 19)
t_bulk = time.time() - t0
print(json.dumps({"record": os.path.basename(record), "pairs": pairs, "features": FEATURES,
                  "convert_game_positions_per_s": round(pairs / t_gen, 1),
                  "bulk_pack_and_encode_positions_per_s": round(len(packed) / t_bulk, 1),
                  "planes_shape": list(planes.shape), "profile": "bench_results.prof"}))
