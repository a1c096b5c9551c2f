#!/bin/bash
# Builds libssn_host.so with AddressSanitizer / ThreadSanitizer and runs the host-side tests (packer, region shards)
# against it.  SURVEY.md §5: the reference gates its integration test on valgrind; this is the multi-threaded
# equivalent for the packer.  -s: a report printed inside a passing test must reach the log.  Usage: tools/sanitize_host.sh address|thread [pytest args]; log -> profiles/r02/.
set -euo pipefail
KIND="${1:-address}"; shift || true
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="$ROOT/somatic_sniper_b200/_build/san_$KIND"
mkdir -p "$OUT" "$ROOT/profiles/r02"
SRC="ssn_pool.c ssn_bgzf.c ssn_bamseek.c ssn_bamread.c ssn_fasta.c ssn_pack.c ssn_shard.c ssn_output.c ssn_inflate.c"
FILES=""
for f in $SRC; do [ -f "$ROOT/somatic_sniper_b200/csrc/host/$f" ] && FILES="$FILES $ROOT/somatic_sniper_b200/csrc/host/$f"; done
gcc -g -O1 -fno-omit-frame-pointer -fsanitize=$KIND -fPIC -shared -std=gnu11 -pthread -I"$ROOT/include" \
    -I"$ROOT/somatic_sniper_b200/csrc/host" $FILES -o "$OUT/libssn_host.so" -lz -lm
LIBSAN=$(gcc -print-file-name=$([ "$KIND" = address ] && echo libasan.so || echo libtsan.so))
LOG="$ROOT/profiles/r02/host_sanitizer_$KIND.log"
cd "$ROOT"
set +e
SSN_SKIP_BUILD=1 SSN_HOST_LIB="$OUT/libssn_host.so" LD_PRELOAD="$LIBSAN" ASAN_OPTIONS=detect_leaks=0:abort_on_error=0 TSAN_OPTIONS="report_signal_unsafe=0 history_size=4" \
    python -m pytest tests/test_shards.py tests/test_packer.py -q -x -s -p no:cacheprovider \
    -k "${SAN_K:-shards or packer_matches or many_small or keep_rule or long_read or first_read or read_window_route or cigar or service or consumers}" "$@" >"$LOG" 2>&1
RC=$?
set -e
echo "exit code $RC" >>"$LOG"
grep -c "ERROR: AddressSanitizer\|WARNING: ThreadSanitizer" "$LOG" | sed 's/^/sanitizer reports: /' | tee -a "$LOG"
tail -5 "$LOG"
exit $RC
