set -x
timeout 1000 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/run_config2_shard.py 2>&1 | tail -7
STORAGE=f32 python tools/run_config2_shard.py 2>&1 | tail -7
