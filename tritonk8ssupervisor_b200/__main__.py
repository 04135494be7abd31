"""``python -m tritonk8ssupervisor_b200 [-c]`` -- drop-in for ``./setup.sh [-c]``
(run it from the workspace root, the directory holding ``terraform/`` and
``ansible/``; setup.sh:9,14,112,154)."""
import sys

from .cli import main

sys.exit(main())
