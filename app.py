"""`uvicorn app:app` — the module name the reference Deployment starts (cluster-config/apps/sd15-api/deployment.yaml:42),
served by b200sd (k8s-nvidia-gpus_b200/server.py).  The pipeline is built on the first /generate."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
app = importlib.import_module("k8s-nvidia-gpus_b200.server").app
