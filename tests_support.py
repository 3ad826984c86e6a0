"""Duck-typed vLLM config / SchedulerOutput builders shared by tests/ and bench.py (no oracle imports here)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests"))
from parity_util import hf_config_from_shape, make_vllm_config, sched_output  # noqa: E402,F401
