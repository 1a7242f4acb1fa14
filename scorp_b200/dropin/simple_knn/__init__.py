"""Import-name shim: `from simple_knn._C import distCUDA2` -> scorp_b200's sm_100a implementation."""
from scorp_b200.simple_knn import _C  # noqa: F401
