#!/usr/bin/env python3
"""HW1 entry (kept under src/ so package_hw.sh picks it up): see b200cnn/runners/hw1.py."""
import sys
from b200cnn.runners.hw1 import main
if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
