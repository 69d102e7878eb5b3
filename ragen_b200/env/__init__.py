"""Environment registry with the reference's plugin point (ragen/env/__init__.py:15-31): classes keyed by `env_type`,
constructed as `cls(config_obj)`.  The classes here are views onto device-resident batches; `BATCH_TYPES` names the
batch class that backs each env type so that EnvStateManager can put all envs of a tag into one batch."""
from .bandit import BanditBatch, BanditEnv, BanditEnvConfig
from .countdown import CountdownBatch, CountdownEnv, CountdownEnvConfig
from .frozen_lake import FrozenLakeBatch, FrozenLakeEnv, FrozenLakeEnvConfig
from .sokoban import SokobanBatch, SokobanEnv, SokobanEnvConfig
from .text_qa import MetaMathQAEnv, MetaMathQAEnvConfig, StaticEnv, StaticEnvConfig, TextScoreBatch

REGISTERED_ENVS = {
    "bandit": BanditEnv,
    "countdown": CountdownEnv,
    "sokoban": SokobanEnv,
    "frozen_lake": FrozenLakeEnv,
    "metamathqa": MetaMathQAEnv,
}

REGISTERED_ENV_CONFIGS = {
    "bandit": BanditEnvConfig,
    "countdown": CountdownEnvConfig,
    "sokoban": SokobanEnvConfig,
    "frozen_lake": FrozenLakeEnvConfig,
    "metamathqa": MetaMathQAEnvConfig,
}

BATCH_TYPES = {
    "bandit": BanditBatch,
    "countdown": CountdownBatch,
    "sokoban": SokobanBatch,
    "frozen_lake": FrozenLakeBatch,
    "metamathqa": TextScoreBatch,
}
