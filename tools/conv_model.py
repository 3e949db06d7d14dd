"""Planning model of the YOLOX-S conv launches (B = 8): FLOPs, MMA-floor cycles, L2->SM operand bytes.
Not part of the product; used to decide tile shapes."""
import sys
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
L = []
def conv(name, cin, cout, k, s, Hin):
    L.append(dict(name=name, cin=cin, cout=cout, k=k, s=s, H=Hin, Ho=Hin // s))
def csp(name, cin, cout, n, H):
    h = cout // 2
    conv(name + ".conv12", cin, 2 * h, 1, 1, H)
    for i in range(n):
        conv(name + ".m%d.conv1" % i, h, h, 1, 1, H)
        conv(name + ".m%d.conv2" % i, h, h, 3, 1, H)
    conv(name + ".conv3", 2 * h, cout, 1, 1, H)
conv("stem", 16, 32, 3, 1, 320)
conv("dark2.0", 32, 64, 3, 2, 320); csp("dark2.1", 64, 64, 1, 160)
conv("dark3.0", 64, 128, 3, 2, 160); csp("dark3.1", 128, 128, 3, 80)
conv("dark4.0", 128, 256, 3, 2, 80); csp("dark4.1", 256, 256, 3, 40)
conv("dark5.0", 256, 512, 3, 2, 40); conv("dark5.1.conv1", 512, 256, 1, 1, 20); conv("dark5.1.conv2", 1024, 512, 1, 1, 20)
csp("dark5.2", 512, 512, 1, 20)
conv("lateral_conv0", 512, 256, 1, 1, 20); csp("C3_p4", 512, 256, 1, 40)
conv("reduce_conv1", 256, 128, 1, 1, 40); csp("C3_p3", 256, 128, 1, 80)
conv("bu_conv2", 128, 128, 3, 2, 80); csp("C3_n3", 256, 256, 1, 40)
conv("bu_conv1", 256, 256, 3, 2, 40); csp("C3_n4", 512, 512, 1, 20)
for l, (C, H) in enumerate(((128, 80), (256, 40), (512, 20))):
    conv("head.stem.%d" % l, C, 128, 1, 1, H); conv("head.conv_a.%d" % l, 128, 256, 3, 1, H)
    conv("head.cls_b.%d" % l, 128, 128, 3, 1, H); conv("head.reg_b.%d" % l, 128, 128, 3, 1, H)
    conv("head.pred.%d" % l, 256, 96, 1, 1, H)
tot_f = tot_cyc = tot_bytes = 0
print("%-22s %5s %5s k s %4s %8s %7s %7s %7s %8s %6s" % ("name", "cin", "cout", "Ho", "GFLOP", "tiles", "mma_us", "l2_us", "hbm_us", "AI"))
rows = []
for c in L:
    M = B * c["Ho"] ** 2
    K = c["cin"] * c["k"] ** 2
    fl = 2.0 * M * K * c["cout"]
    BN = min(128, c["cout"]) if c["cout"] % 128 == 0 or c["cout"] < 128 else 96
    ntn = c["cout"] // BN if c["cout"] % BN == 0 else 1
    mt = (M + 127) // 128
    tiles = mt * ntn
    # mma cycles per (128 x BN x 16): measured issue cost 45/47/61/120 for N=32/64/128/256
    cyc16 = {32: 45, 64: 47, 96: 55, 128: 64}.get(BN, 64)
    mma_cyc = tiles * (K // 16) * cyc16 / 148.0
    # L2->SM bytes, current plan: per tile per (tap, kchunk): A 128*ck*2 + B BN*ck*2
    a_bytes = tiles * K * 128 * 2
    b_bytes = tiles * K * BN * 2
    l2 = a_bytes + b_bytes
    hbm = M * c["cout"] * 2 + B * c["H"] ** 2 * c["cin"] * 2 + K * c["cout"] * 2
    rows.append((c, fl, tiles, mma_cyc, l2, hbm))
    tot_f += fl; tot_cyc += mma_cyc; tot_bytes += l2
    print("%-22s %5d %5d %d %d %4d %8.2f %7d %7.1f %7.1f %7.1f %6.0f" % (c["name"], c["cin"], c["cout"], c["k"], c["s"], c["Ho"], fl / 1e9, tiles,
          mma_cyc / 1965.0, l2 / 12.4e6, hbm / 6.5e6, fl / l2))
print("total GFLOP %.1f  mma floor %.0f us  L2->SM %.2f GB = %.0f us at 12.4 TB/s" % (tot_f / 1e9, tot_cyc / 1965.0, tot_bytes / 1e9, tot_bytes / 12.4e6))
