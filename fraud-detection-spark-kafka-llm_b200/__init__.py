"""B200-native drop-in for the Spark ML PipelineModel hot path of
wangwang2111/fraud-detection-spark-kafka-llm (dialogue text -> TF-IDF -> LR / DT / RF / GBT), plus
random-forest histogram training.  The directory name is not a Python identifier: import it with
``importlib.import_module("fraud-detection-spark-kafka-llm_b200")`` or through the ``fdb200`` alias
module at the repository root."""
from . import spark_format
from ._lib import FdError, Handle, LIB_PATH, load_library
from .evaluation import BinaryClassificationEvaluator, MulticlassClassificationEvaluator, confusion_matrix
from .hostmem import gpu_numa_node, numa_local
from .packing import PackedText, pack_strings, shard_bounds
from .pipeline import Pipeline, PipelineModel, Row, TransformResult, predict_and_get_label
from .serving import MicroBatcher, classify_batch
from .stages import (CountVectorizer, CountVectorizerModel, DecisionTreeClassificationModel, DecisionTreeClassifier, RandomForestClassifier, GBTClassificationModel, HashingTF, IDF, IDFModel,
                     LogisticRegressionModel, RandomForestClassificationModel, StopWordsRemover, Tokenizer,
                     load_default_stop_words, load_stage)

__all__ = [n for n in dir() if not n.startswith("_")]
