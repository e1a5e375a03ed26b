for c in 128 192 256; do echo "chunk $c: $(SMIL_BWD_CHUNK=$c python tools_quick_bench.py mouse 1024 2>&1 | tail -1)"; done
