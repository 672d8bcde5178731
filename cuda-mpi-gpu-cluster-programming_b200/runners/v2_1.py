"""V2.1 entry point (python -m b200cnn.runners.v2_1); see runners/v2.py."""
import sys
from .v2 import main_v21
if __name__ == "__main__":
    sys.exit(main_v21(sys.argv[1:]))
