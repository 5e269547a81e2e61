"""Drop-in namespace: the upstream package's import name (`cnmf`,
/root/reference/setup.py:9) re-exporting the B200-native entry point."""
from cnmf_b200 import NMF, Comm, Plan, load_result, nmf, nmf_restarts, save_result  # noqa: F401
