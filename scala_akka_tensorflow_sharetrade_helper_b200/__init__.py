"""B200-native Q-learning decision policy: the one hot path of
cosmir17/Scala-akka-tensorflow-sharetrade-helper, as sm_100a CUDA kernels behind a C ABI.

Layout: csrc/ (kernels + C ABI implementation), _build (nvcc, in-tree), _capi (ctypes stub of
include/qpolicy.h), policy / trainer (host-side mirrors of the reference's actor messages).
"""
from ._build import build, LIB_PATH  # noqa: F401
from .ingest import parse_prices, load_price_file  # noqa: F401
from .policy import (Action, BatchingPolicy, Buy, Hold, IllegalArgumentException, PolicyConfig, QDecisionPolicy,  # noqa: F401
                     QPolicyError, SelectionAction, Sell, UpdateQ, Updated)
from .sharding import shard  # noqa: F401
from .trainer import (Initialised, NotComputed, StockDataResponse, Trained, TrainedData,  # noqa: F401
                      TrainerChild, TrainerRouter)
