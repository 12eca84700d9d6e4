"""Stand-in for ssh in the launcher test: `fake_ssh.py <host> <command>` runs the command locally."""
import subprocess
import sys

sys.exit(subprocess.call(["sh", "-c", sys.argv[2]]))
