"""Alias so that ``import fdb200`` gives the package in ``fraud-detection-spark-kafka-llm_b200/``."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
sys.modules[__name__] = importlib.import_module("fraud-detection-spark-kafka-llm_b200")
