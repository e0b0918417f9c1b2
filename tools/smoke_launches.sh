#!/bin/bash
# Launch list of __graft_entry__.smoke() with device times (what the driver records at round end).
set -e
cd "$(dirname "$0")/.."
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_smoke.csv python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_ncu.log 2>&1
python - <<'PY'
import csv
rows = [r for r in csv.reader(open("gpurun_out/r02_launches_smoke.csv")) if len(r) > 10 and r[0].isdigit()]
for r in rows:
    print(r[4][:60].ljust(62), r[-1], r[-2])
PY
