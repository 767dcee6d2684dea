import subprocess, sys
base = [sys.executable, "tools/vs_one_rank.py", "--rank", "0", "--world", "2", "--size", "1024", "--dtype", "f32", "--slices", "1", "--frames", "1", "--first", "1"]
def fails(r, m, tile=64):
    p = subprocess.run(base + ["--part", str(r), str(m), "--tile", str(tile)], capture_output=True, text=True, timeout=200)
    return p.returncode != 0
x0, y0, size = 704, 960, 64
for T in (32, 16, 8):
    n = 2048 // T
    found = None
    for dy in (0, 1):
        for dx in (0, 1):
            tx, ty = (x0 + dx * T) // T, (y0 + dy * T) // T
            f = fails(ty * n + tx, n * n, T)
            print("tile", T, "at", tx * T, ty * T, "fails" if f else "ok", flush=True)
            if f and found is None:
                found = (tx * T, ty * T)
    if found is None:
        print("none fails alone")
        break
    x0, y0 = found
print("box", x0, y0)
