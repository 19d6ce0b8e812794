"""``python -m fedb200.client -a host:port [-c Y] [-r] [--lr LR] ...``
(reference: ``python3 client.py``, src/client.py:55-71)."""
import sys

from .control.cli import client_main

if __name__ == "__main__":
    sys.exit(client_main())
