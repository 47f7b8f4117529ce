python -c "
import sys; sys.path.insert(0, '.'); import bench
cfg, files, d, _ = bench.fastq_files(0, 1000000, device='cuda'); gz, _ = bench.gzip_copies(files, d); print(' '.join(gz))
" > /tmp/files.txt
F=$(tail -1 /tmp/files.txt)
for i in 1 2 3 4 5 6 7 8; do rm -rf /tmp/o; mkdir /tmp/o; s=$(date +%s.%N); HUMID_TIMING=1 ./humid_b200/humid -n 24 -m 1 -x -s -d /tmp/o -l /tmp/o/log $F 2> /tmp/err.txt; e=$(date +%s.%N); echo "run $i total $(python3 -c "print(round($e - $s, 3))")"; grep -E "create|Reading|locked|Writing|destroy" /tmp/err.txt | tr '\n' ';'; echo; done
