#!/usr/bin/env bash
# Build the sm_100a library, run the test-suite (CPU part always, GPU part when a device is visible), write a JUnit
# report and pack a source archive.  Counterpart of the reference's test_and_make_submission.sh (pytest + zip).
set -euo pipefail
cd "$(dirname "$0")/.."
python -c "import __graft_entry__ as g; g.build()"
mkdir -p build/reports
python -m pytest tests -q -m "not gpu" --junitxml=build/reports/cpu.xml
if python -c "import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)"; then
  python -m pytest tests -q -m gpu --junitxml=build/reports/gpu.xml
fi
git archive --format=zip -o build/cs336_a2_systems_b200-src.zip HEAD
echo "wrote build/reports/*.xml and build/cs336_a2_systems_b200-src.zip"
