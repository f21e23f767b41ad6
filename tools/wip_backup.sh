#!/bin/bash
# Snapshot the whole working tree (tracked + untracked, minus ignored) as a commit on branch wip-backup without
# touching the current branch, index or files.  Protection against losing uncommitted work.
set -e
cd "$(git rev-parse --show-toplevel)"
export GIT_INDEX_FILE=$(mktemp)
cp .git/index "$GIT_INDEX_FILE"
git add -A
tree=$(git write-tree)
parent=$(git rev-parse HEAD)
c=$(git commit-tree "$tree" -p "$parent" -m "wip backup $(date -u +%H:%M:%S): ${1:-work in progress}")
git branch -f wip-backup "$c"
rm -f "$GIT_INDEX_FILE"
echo "wip-backup -> $c"
