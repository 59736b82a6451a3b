"""Runs every GPU test in its own process (a faulting kernel poisons the CUDA context of the whole process, so
one crash would otherwise fail every later test).  Usage: python tests/run_isolated.py [-j N] [pytest args]"""
import concurrent.futures as cf
import subprocess
import sys


def main():
    args = sys.argv[1:]
    jobs = 4
    if args[:1] == ["-j"]:
        jobs, args = int(args[1]), args[2:]
    sel = args or ["tests", "-m", "gpu"]
    out = subprocess.run([sys.executable, "-m", "pytest", "--collect-only", "-q"] + sel, capture_output=True, text=True)
    ids = [l.strip() for l in out.stdout.splitlines() if "::" in l]
    print(f"{len(ids)} tests", flush=True)

    def run(tid):
        r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "--timeout", "600", "-p", "no:cacheprovider", tid],
                           capture_output=True, text=True)
        return tid, r.returncode, r.stdout[-3500:] + r.stderr[-1500:]

    fails = 0
    with cf.ThreadPoolExecutor(jobs) as ex:
        for tid, rc, log in ex.map(run, ids):
            print(("PASS " if rc == 0 else "FAIL ") + tid, flush=True)
            if rc != 0:
                fails += 1
                print(log, flush=True)
    print(f"== {len(ids) - fails} passed, {fails} failed")
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
