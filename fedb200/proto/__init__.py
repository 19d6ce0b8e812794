"""federated.proto contract: messages, stub, servicer (see contract.py)."""
from .contract import *  # noqa: F401,F403
from .contract import (DESCRIPTOR, TrainerStub, TrainerServicer, add_TrainerServicer_to_server,
                       make_channel, make_server, file_descriptor_proto, method_path)
