"""Writes tests/golden/dlrm_tablewise_26x8.strategy with the REFERENCE's own writer
(save_strategies_to_file, /root/reference/src/runtime/strategy.cc:147-189, compiled unmodified into
oracle/_ref/libffref_cpu.so): the DLRM generator's table-wise strategy for 26 tables on 8 GPUs
(examples/cpp/DLRM/strategies/dlrm_strategy.cc:252-262).  Run in the build container:
    python tests/golden/make_strategy_golden.py"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle as O  # noqa: E402

if __name__ == "__main__":
    O.ref_strategy_save_tablewise(os.path.join(HERE, "dlrm_tablewise_26x8.strategy"), 26, 8)
    print(open(os.path.join(HERE, "dlrm_tablewise_26x8.strategy")).read()[:200])
