#!/usr/bin/env python3
"""Root-level entry point kept for CLI compatibility with the reference's `log_analysis.py`
(ingest / query / stats / speedup / plot / export).  Implementation: b200cnn/harness/log_analysis.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from b200cnn.harness.log_analysis import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
