"""fedb200 — a Blackwell-native federated-learning framework (import name: ``fedb200``).

The directory is called ``federated-learning-with-mpi_b200`` (not an identifier); the sibling
``fedb200/`` package points its ``__path__`` here, so ``import fedb200`` /
``from fedb200.fed.engine import FedEngine`` are the supported spellings.
"""
from .data.income import Prepared, load_table, prepare, synthetic_income_like  # noqa: F401
from .data.partition import make_shards  # noqa: F401
from .fed.engine import FedEngine  # noqa: F401
from .fed.problem import FedProblem, TrainHyper  # noqa: F401
from .models.spec import MLPSpec  # noqa: F401

__version__ = "0.1.0"
