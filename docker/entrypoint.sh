#!/bin/sh
# Starts sshd when asked to (the `ssh` launcher of DriverSession logs into this container), then runs the command.
set -e
if [ "${METISFL_SSHD:-0}" = "1" ]; then
    ssh-keygen -A >/dev/null
    /usr/sbin/sshd
fi
exec "$@"
