"""Environment plugin interface, same names and call contract as ragen/env/base.py:5-100.

An env is constructed as `cls(config)`, and the manager calls `reset(seed=..., mode=...)`,
`step(action) -> (obs, reward, done, info)`, `render()`, `close()`; discrete envs also expose
`get_all_actions()`.  Same seed => same env.
"""
from abc import ABC, abstractmethod
from typing import Any, Dict, List, Tuple


class BaseEnv(ABC):
    @abstractmethod
    def reset(self, seed=None, **kwargs) -> Any:
        ...

    @abstractmethod
    def step(self, action) -> Tuple[Any, float, bool, Dict]:
        ...

    def render(self, mode: str = "text") -> Any:
        return None

    def compute_reward(self, action, **kwargs) -> float:
        return 0.0

    def close(self):
        return None


class BaseDiscreteActionEnv(BaseEnv, ABC):
    @abstractmethod
    def get_all_actions(self) -> List[int]:
        ...


class BaseLanguageBasedEnv(BaseEnv, ABC):
    pass


class BaseEnvConfig(ABC):
    def __init__(self):
        self.invalid_act = ""
        self.invalid_act_score = 0
