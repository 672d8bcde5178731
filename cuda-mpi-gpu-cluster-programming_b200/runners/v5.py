"""V5 entry point (python -m b200cnn.runners.v5); see runners/v4.py."""
import sys
from .v4 import main_v5
if __name__ == "__main__":
    sys.exit(main_v5(sys.argv[1:]))
