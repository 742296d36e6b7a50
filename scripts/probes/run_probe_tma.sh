# every case in its own process (a rejected coordinate kills the CUDA context)
for x in 0 16 32 48 4 8 20 24 1 2 3 17 47 601 624 630; do timeout 60 ./scripts/probes/probe_tma_u8 $x 16 0 | head -1; done
for y in 0 1 17 47 450 470; do timeout 60 ./scripts/probes/probe_tma_u8 32 $y 0 | head -1; done
timeout 60 ./scripts/probes/probe_tma_u8 48 47 1 | head -1
timeout 60 ./scripts/probes/probe_tma_u8 48 47 2 | head -1
