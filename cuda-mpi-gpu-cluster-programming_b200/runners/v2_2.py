"""V2.2 entry point (python -m b200cnn.runners.v2_2); see runners/v2.py."""
import sys
from .v2 import main_v22
if __name__ == "__main__":
    sys.exit(main_v22(sys.argv[1:]))
