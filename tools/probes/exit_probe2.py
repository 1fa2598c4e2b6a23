import subprocess
import sys

r = subprocess.run([sys.executable, "-X", "faulthandler", "tools/lat_probe.py"], capture_output=True, text=True)
print("rc", r.returncode)
print(r.stdout[-600:])
print(r.stderr[-3000:])
