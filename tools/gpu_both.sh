python -m pytest tests -x -q -m gpu 2>&1 | grep -E "^E  |passed|failed|Error|^FAILED" | head -12
CFGS="${CFGS:-C1 C2 C3 C4 C5 W1080 W720 C3P}" bash tools/gpu_allcfg.sh
