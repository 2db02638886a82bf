set -x
python -m pytest tests -m gpu -x -q -k "loss" 2>&1 | tail -25
