"""Job flags and defaults.

Reference: `--key value` pairs parsed by Flink's ParameterTool (Job.scala:114) with the defaults
of utils/DefaultJobParameters.scala:4-11 and the flag table in SURVEY §5.6.  Kafka-address flags
are accepted for command-line compatibility and ignored (topics are in-process ring buffers).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

DEFAULT_JOB_NAME = "OML_job_1"            # DefaultJobParameters.scala:4
DEFAULT_PARALLELISM = 16                  # :5
DEFAULT_MAX_MSG_PARAMS = 2000             # :6
DEFAULT_TEST = True                       # :9
DEFAULT_TIMEOUT_MS = 30000                # :10
DEFAULT_TEST_SET_SIZE = 256               # :11
RECORD_BUFFER_CAP = 100000                # SpokeLogic.scala:32
REQUEST_BUFFER_CAP = 10000                # SpokeLogic.scala:35
HUB_CACHE_CAP = 20000                     # StateAccumulators.scala:131
TICK_EVERY = 100                          # FlinkSpoke.scala:85
TOPICS = ("trainingData", "forecastingData", "requests", "responses", "predictions", "psMessages", "performance")


def _to_bool(v) -> bool:
    return str(v).lower() in ("1", "true", "yes", "on")


@dataclass
class JobParameters:
    jobName: str = DEFAULT_JOB_NAME
    parallelism: int = DEFAULT_PARALLELISM
    maxMsgParams: int = DEFAULT_MAX_MSG_PARAMS
    test: bool = DEFAULT_TEST
    timeout: int = DEFAULT_TIMEOUT_MS
    testSetSize: int = DEFAULT_TEST_SET_SIZE
    checkpointing: bool = False
    local: bool = True
    checkInterval: int = 5000
    stateBackend: str = "checkpoints"
    topics: Dict[str, str] = field(default_factory=lambda: {t: t for t in TOPICS})
    addrs: Dict[str, str] = field(default_factory=dict)
    # extensions of this build
    engine: str = "gpu"            # gpu | oracle
    distributedKMeans: bool = True
    injectStall: Optional[str] = None   # "rank:ms" fault-injection hook
    extra: Dict[str, str] = field(default_factory=dict)

    @classmethod
    def from_args(cls, argv: List[str]) -> "JobParameters":
        p = cls()
        i = 0
        while i < len(argv):
            a = argv[i]
            if not a.startswith("--"):
                i += 1
                continue
            key = a[2:]
            if "=" in key:
                key, val = key.split("=", 1)
            elif i + 1 < len(argv) and not argv[i + 1].startswith("--"):
                val = argv[i + 1]
                i += 1
            else:
                val = "true"
            p.set(key, val)
            i += 1
        return p

    def set(self, key: str, val: str) -> None:
        if key.endswith("Topic") and key[:-5] in TOPICS:
            self.topics[key[:-5]] = val
        elif key.endswith("Addr"):
            self.addrs[key[:-4]] = val
        elif key in ("parallelism", "maxMsgParams", "timeout", "testSetSize", "checkInterval"):
            setattr(self, key, int(val))
        elif key in ("test", "checkpointing", "local", "distributedKMeans"):
            setattr(self, key, _to_bool(val))
        elif key in ("jobName", "stateBackend", "engine", "injectStall"):
            setattr(self, key, val)
        else:
            self.extra[key] = val

    def get(self, key: str, default=None):
        return getattr(self, key, self.extra.get(key, default))
