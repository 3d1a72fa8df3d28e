"""Synthetic raw-event files for the featurisation tests (flow_post_lda.scala /
dns_post_lda.scala inputs).  Small, seeded, and salted with the rows the
scripts' filters and string rules care about."""
import os
import random


def _probs(rnd, k=20):
    v = [rnd.random() ** 3 + 1e-6 for _ in range(k)]
    s = sum(v)
    return " ".join("%.12g" % (x / s) for x in v)


FLOW_HEADER = ("tr,try,trm,trd,trh,trm,trs,td,sa,da,sp,dp,pr,flg,fwd,stos,ipkt,ibyt,opkt,obyt,in,out,sas,das,"
               "dtos,dir,ra")


def flow_row(rnd, ips):
    h, m, s = rnd.randint(0, 23), rnd.randint(0, 59), rnd.randint(0, 59)
    sp = rnd.choice([0, 22, 53, 80, 443, 1024, 1025, 8080, rnd.randint(1, 65535), rnd.randint(1025, 65535)])
    dp = rnd.choice([0, 25, 80, 443, 1024, 1025, 3389, rnd.randint(1, 65535), rnd.randint(1025, 65535)])
    ipkt = rnd.choice([1, 1, 2, 3, 5, 8, rnd.randint(1, 2000)])
    ibyt = ipkt * rnd.choice([40, 64, 120, 576, 1500, rnd.randint(40, 1500)])
    return ["2015-04-12 %02d:%02d:%02d" % (h, m, s), "2015", "4", "12", str(h), str(m), str(s),
            "%.3f" % (rnd.random() * 30), rnd.choice(ips), rnd.choice(ips), str(sp), str(dp),
            rnd.choice(["TCP", "UDP", "ICMP"]), ".AP.SF", "0", "0", str(ipkt), str(ibyt), "0", "0", "2", "3", "0",
            "0", "0", "0", "10.219.32.250"]


def make_flow(dirname, n=3000, seed=7):
    """Writes flow.csv, doc_results.csv, word_results.csv under dirname; returns their paths."""
    rnd = random.Random(seed)
    ips = ["10.%d.%d.%d" % (rnd.randint(0, 2), rnd.randint(0, 255), rnd.randint(0, 255)) for _ in range(150)]
    ips += ["9.9.9.9", "192.168.1.33"]
    lines = [FLOW_HEADER]
    for i in range(n):
        f = flow_row(rnd, ips)
        kind = rnd.random()
        if kind < 0.01:
            lines.append(",".join(f[:26]))            # short row: dropped
        elif kind < 0.02:
            lines.append(",".join(f + ["extra"]))     # long row: dropped
        elif kind < 0.03:
            lines.append(",".join(f[:26]) + ",")      # trailing empty field vanishes in split: dropped
        elif kind < 0.04:
            lines.append(FLOW_HEADER)                 # every copy of the header goes
        elif kind < 0.06:
            lines.append("  " + ",".join(f) + " \t")  # trimmed before the second split
        elif kind < 0.07:
            f[4] = " %s " % f[4]                       # toDouble tolerates blanks
            lines.append(",".join(f))
        elif kind < 0.08:
            f[17] = "1.5e3"
            lines.append(",".join(f))
        else:
            lines.append(",".join(f))
    flow = os.path.join(dirname, "flow.csv")
    with open(flow, "w") as fh:
        fh.write("\n".join(lines) + "\n")
    # topic mixes for most (not all) IPs; one repeated key: the later line wins
    docs = os.path.join(dirname, "doc_results.csv")
    with open(docs, "w") as fh:
        for ip in ips[:120]:
            fh.write("%s,%s\n" % (ip, _probs(rnd)))
        fh.write("%s,%s\n" % (ips[3], _probs(rnd)))
        fh.write('%s,"%s"\n' % ("192.168.1.33", _probs(rnd)))  # quotes are stripped by the script
    return flow, docs, os.path.join(dirname, "word_results.csv")


def write_word_results(path, words, seed=11, keep=0.8, cut20=True):
    """word_results.csv for the given word strings: names cut to 20 bytes as lda_post.py's S20
    array does, a fraction left out so that lookups miss."""
    rnd = random.Random(seed)
    seen = set()
    with open(path, "w") as fh:
        for w in words:
            name = w[:20] if cut20 else w
            if name in seen or rnd.random() > keep:
                continue
            seen.add(name)
            fh.write("%s,%s\n" % (name, _probs(rnd)))


DNS_NAMES = ["www.google.com", "mail.google.com", "a.b.c.example.co.uk", "intel.com", "download.intel.com",
             "x1.y2.intel.com", "10.104.103.123.in-addr.arpa", "localhost", "example.org", "cdn.assets.shop.de",
             "weird..dots.net", "trailing.dot.com.", "", "...", "a.very-long-subdomain-name-1234567890.tracker.io",
             "ns1.krd", "foo.bar.krd", "q.w.e.r.t.y.ru"]


def make_dns(dirname, n=3000, seed=13):
    rnd = random.Random(seed)
    ips = ["172.16.%d.%d" % (rnd.randint(0, 3), rnd.randint(0, 255)) for _ in range(100)]
    lines = []
    t0 = 1442171291
    for i in range(n):
        name = rnd.choice(DNS_NAMES)
        if rnd.random() < 0.3:
            name = "".join(rnd.choice("abcdefghijklmnopqrstuvwxyz0123456789-") for _ in range(rnd.randint(1, 30))) + \
                   "." + name
        f = ["Sep 13 2015 19:08:%02d UTC" % rnd.randint(0, 59), str(t0 + rnd.randint(0, 86399)),
             str(rnd.choice([60, 74, 90, 120, 300, rnd.randint(60, 1500)])), rnd.choice(ips), name, "0x00000001",
             str(rnd.choice([1, 1, 1, 28, 12, 15, 16])), str(rnd.choice([0, 0, 0, 3, 2]))]
        kind = rnd.random()
        if kind < 0.01:
            lines.append(",".join(f[:7]))
        elif kind < 0.02:
            lines.append("Sep 13, 2015 19:08:11 UTC," + ",".join(f[1:]))  # the comma adds a field: dropped
        else:
            lines.append(",".join(f))
    dns = os.path.join(dirname, "dns.csv")
    with open(dns, "w") as fh:
        fh.write("\n".join(lines) + "\n")
    top = os.path.join(dirname, "top-1m.csv")
    with open(top, "w") as fh:
        for i, d in enumerate(["google.com", "example.org", "shop.de", "tracker.io", "yandex.ru", "dots.net"]):
            fh.write("%d,%s\n" % (i + 1, d))
    docs = os.path.join(dirname, "doc_results.csv")
    with open(docs, "w") as fh:
        for ip in ips[:80]:
            fh.write("%s,%s\n" % (ip, _probs(rnd)))
    return dns, top, docs, os.path.join(dirname, "word_results.csv")
