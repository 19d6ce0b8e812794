"""Flag surface: every reference flag keeps its spelling and default; what the reference hard-codes is a flag.
reference: src/server.py:270-279, src/client.py:56-69, src/main.py:20-28."""
import subprocess
import sys

from fedb200 import main as M
from fedb200.control import cli


def test_server_flags_of_the_reference():
    a = cli.server_parser().parse_args([])
    assert (a.compressFlag, a.p, a.backupAddress, a.backupPort) == (None, "n", "localhost", "8080")
    a = cli.server_parser().parse_args(["-c", "Y", "--p", "y", "--backupAddress", "10.0.0.7", "--backupPort", "9000"])
    assert (a.compressFlag, a.p, a.backupAddress, a.backupPort) == ("Y", "y", "10.0.0.7", "9000")
    # hard-coded in the reference (B10): worker list, 20 rounds, MobileNet, 10 s watchdog
    assert a.workers == "localhost:50051,localhost:50052" and a.rounds == 20 and a.model == "MobileNet" and a.watchdog == 10.0
    b = cli.server_parser().parse_args(["--workers", "h1:1,h2:2,h3:3", "--rounds", "5", "--plane", "peer", "--weighted"])
    assert b.workers.split(",") == ["h1:1", "h2:2", "h3:3"] and b.rounds == 5 and b.plane == "peer" and b.weighted


def test_client_flags_of_the_reference():
    a = cli.client_parser().parse_args([])
    assert (a.compressFlag, a.address, a.resume, a.lr) == (None, "temp", False, 0.1)
    a = cli.client_parser().parse_args(["--compressFlag", "Y", "-a", "localhost:50051", "-r", "--lr", "0.01"])
    assert (a.compressFlag, a.address, a.resume, a.lr) == ("Y", "localhost:50051", True, 0.01)
    assert a.plane == "legacy" and a.codec == "none" and a.batch_size == 128 and a.data_root is None


def test_main_flags_of_the_reference():
    a = M.parser().parse_args([])
    assert a.lr == 0.1 and a.resume is False and a.model == "MobileNet" and a.t_max == 200
    a = M.parser().parse_args(["--lr", "0.02", "--resume"])
    assert a.lr == 0.02 and a.resume is True


def test_module_entry_points_answer_help():
    for mod in ("fedb200.server", "fedb200.client", "fedb200.main"):
        out = subprocess.run([sys.executable, "-m", mod, "--help"], capture_output=True, text=True, timeout=120)
        assert out.returncode == 0 and "usage" in out.stdout.lower(), (mod, out.stderr[-300:])
