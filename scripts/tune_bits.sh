#!/bin/bash
# chain geometry sweep for the moment tree (env CF_TREE_BITS, csrc/tree.cu); "auto" = the launcher's choice
run() { python scripts/time_tree.py $1 $2 $3 | grep tree | sed 's/launches.*->/->/'; }
for b in auto 8,8 9,8 9,7 6,5,5 7,9 16 ; do
  echo "== bits $b"; if [ $b = auto ]; then unset CF_TREE_BITS; else export CF_TREE_BITS=$b; fi
  run 1e8 0.05 3
done
for b in auto 9 5,4 8 6,4 10; do
  echo "== bits $b"; if [ $b = auto ]; then unset CF_TREE_BITS; else export CF_TREE_BITS=$b; fi
  run 1e6 0.05 3
done
for b in auto 7,7 8,6 9,5 5,5,4; do
  echo "== bits $b"; if [ $b = auto ]; then unset CF_TREE_BITS; else export CF_TREE_BITS=$b; fi
  run 2e6 0.8 5
done
for b in auto 6 5 4 3,3; do
  echo "== bits $b"; if [ $b = auto ]; then unset CF_TREE_BITS; else export CF_TREE_BITS=$b; fi
  run 16384 0.3 2
done
