timeout 900 python tools/gn_scale.py rays:64 calib:64 points:256 2>&1 | tail -5
