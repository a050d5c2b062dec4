"""Reads the POZ_KERNEL_TRACE dump of scripts/micro/step_probe.py (stdin or a file) and prints, per traced step,
the window between the end of fwd_l1 and the start of bwd_x_l1 (= the head kernel plus two launch gaps) and the
window between the end of bwd_w_l0 and the start of the next step's fwd_l0 (= the update kernel plus gaps)."""
import re, sys
txt = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
items = re.findall(r"first CTA entry at (-?\d+) ns, last CTA exit at (-?\d+) ns[^\n]*\n\[poz kernel trace\] tag (\d+) M (\d+) N (\d+) K (\d+)", txt)
items = [(int(a), int(b), int(t), int(M), int(N), int(K)) for a, b, t, M, N, K in items][::-1]   # latest first
head, upd = [], []
for i, (a, b, tag, M, N, K) in enumerate(items):          # printed latest first
    if tag & 1 and not tag & 2 and N == 128:               # fwd_l1
        later = []
        for j in range(i - 1, -1, -1):
            if items[j][2] & 3 == 1 and items[j][4] == 256: break      # fwd_l0 of the next step
            if items[j][2] & 3 == 3: later.append(items[j][0])          # fragments of bwd_x_l1
        if later: head.append(min(later) - b)
    if tag & 3 == 1 and N == 256:                          # fwd_l0: the previous step's bwd_w_l0 is printed after it
        ends = []
        for j in range(i + 1, len(items)):
            if items[j][2] & 3 == 1 and items[j][4] == 256: break
            if items[j][2] & 4 and items[j][3] == 784: ends.append(items[j][1])
        if ends: upd.append(a - max(ends))
print("head window ns:", head, " update window ns:", upd)
