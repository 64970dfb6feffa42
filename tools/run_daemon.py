#!/usr/bin/env python3
"""Runs the device plugin daemon (k8s-gpu-sharing-plugin_b200/lifecycle.py) on this node.
   python tools/run_daemon.py --resource-config gpu:shared-gpu:4 [--plugin-dir DIR] [--replica-topology]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

entry.load_package()
from k8s_gpu_sharing_plugin_b200 import lifecycle  # noqa: E402

if __name__ == "__main__":
    sys.exit(lifecycle.main())
