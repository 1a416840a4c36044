python tools/small_host_profile.py 2>&1 | tail -4
