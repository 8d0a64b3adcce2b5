"""Minimal stand-in for the third-party `kubernetes` client (TEST INFRASTRUCTURE).

The reference imports `kubernetes` at module top (/root/reference/check-gpu-node.py:34-35)
and the real package (pinned 33.1.0, /root/reference/uv.lock:104-105) is not installable
here (no network).  Only the names the reference touches exist: `client.{V1Node,
V1NodeCondition, CoreV1Api}` and `config.load_kube_config`, plus the plain model classes
needed to build fixtures with the attribute names the reference reads.
"""
from . import client, config  # noqa: F401

__version__ = "33.1.0+stub"
