for ws in 13 8 1; do for args in "16384 1 1" "16384 0 1" "16384 1 3" "400 0 1" "400 1 2" "40000 1 2"; do
  echo "ws=$ws args=$args: $(PGA_WS=$ws timeout 60 python scripts/dbg_ws.py $args 2>&1 | tail -1)"
done; done
