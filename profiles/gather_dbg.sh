for nt in 0 1; do
for cfg in "4 2 1024" "4 2 2048" "6 2 2048" "8 2 1024"; do
  set -- $cfg
  for rep in 1 2; do
    echo "== nt=$nt T=$1 S=$2 C=$3 rep $rep: $(CALIB_GATHER_NT=$nt CALIB_GATHER_THREADS=$1 CALIB_GATHER_SLOTS=$2 CALIB_GATHER_CHUNK_KB=$3 python profiles/gather_sweep.py one)"
  done
  CALIB_GATHER_NT=$nt CALIB_DEBUG_TIMING=1 CALIB_GATHER_THREADS=$1 CALIB_GATHER_SLOTS=$2 CALIB_GATHER_CHUNK_KB=$3 python profiles/gather_sweep.py one 2>&1 | grep -E "gather thread|upload_views|calib_create" | tail -$(( $1 + 2 ))
done
done
