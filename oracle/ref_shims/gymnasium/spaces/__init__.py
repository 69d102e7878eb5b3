from . import discrete
from .discrete import Discrete
