"""`load_kube_config` is a recorded no-op: there is no cluster (check-gpu-node.py:160-169)."""
calls = []


def load_kube_config(config_file=None, **kw):
    calls.append(config_file)
