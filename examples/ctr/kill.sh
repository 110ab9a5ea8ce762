#!/bin/bash
# stop the local PS roles started by a launch script (by process group of this shell, never by name pattern)
kill -- -$$ 2>/dev/null
