run() { python bench.py --no-extras --sustain-s 0 --steps 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1', d['value'], d['e2e']['value'])
"; }
run base
TFS_X_NOPIN=1 run nopin
TFS_X_FRESH=1 run fresh
TFS_X_NOPIN=1 TFS_X_FRESH=1 run both
