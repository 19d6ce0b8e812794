"""``python -m fedb200.server [--p y] [-c Y] [--backupAddress H] [--backupPort P] ...``
(reference: ``python3 server.py``, src/server.py:268-301)."""
import sys

from .control.cli import server_main

if __name__ == "__main__":
    sys.exit(server_main())
