"""Fixture: the 342-row pair list (trg, src ids only) of the reference's shipped result file
Registration/registration_dataset/reg_result.txt:2-343 (main.py:186-192 reads exactly these two columns).
    python tests/golden/make_reg_pairs.py   ->  tests/golden/reg_pairs.txt
The poses in that file came out of a random RANSAC initialisation and are not reproducible; only
the pair ids are kept."""
import os
SRC = "/root/reference/Registration/registration_dataset/reg_result.txt"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reg_pairs.txt")
with open(SRC) as f:
    lines = f.readlines()
with open(OUT, "w") as f:
    f.write("idx1,idx2\n")
    for ln in lines[1:]:
        a = ln.strip().split(",")
        if len(a) >= 2:
            f.write("%d,%d\n" % (int(float(a[0])), int(float(a[1]))))
print(OUT, sum(1 for _ in open(OUT)) - 1, "pairs")
