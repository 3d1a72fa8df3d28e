"""CPU restatement of flow_post_lda.scala and dns_post_lda.scala, line by line,
in plain Python.  TEST INFRASTRUCTURE ONLY: used by tests/ as the checker of
oni_flow_events_* / oni_dns_events_* (csrc/events.cu); nothing in the product
imports it.

PARITY UNPINNED: the reference scripts need Spark, Scala and a JVM, none of
which exist here or on the GPU box, and the reference holds no
test or fixture for them.  What the scripts leave to the JVM is fixed here the
same way the product fixes it: Double.toString = the shortest decimal that
round-trips (Python's repr digits) in Java's layout; sortByKey ties keep input
order; the entropy terms are added in ascending character order.

Each function names the script lines it follows.
"""
import math
from decimal import Decimal


# ---------------------------------------------------------------- JVM text rules
def jsplit(s, sep):
    """String.split(sep) for a literal separator: trailing empty strings removed."""
    parts = s.split(sep)
    while parts and parts[-1] == "":
        parts.pop()
    if not parts:
        return [""] if s == "" else []
    return parts


def jtrim(s):
    """String.trim: code units <= U+0020 off both ends."""
    a, b = 0, len(s)
    while a < b and ord(s[a]) <= 0x20:
        a += 1
    while b > a and ord(s[b - 1]) <= 0x20:
        b -= 1
    return s[a:b]


def jdouble(v):
    """Double.toString."""
    if v != v:
        return "NaN"
    if v in (float("inf"), float("-inf")):
        return "Infinity" if v > 0 else "-Infinity"
    if v == 0:
        return "-0.0" if math.copysign(1.0, v) < 0 else "0.0"
    sign, digits, exp = Decimal(repr(abs(v))).as_tuple()
    digits = list(digits)
    while len(digits) > 1 and digits[-1] == 0:
        digits.pop()
        exp += 1
    ds = "".join(map(str, digits))
    e10 = len(ds) - 1 + exp  # value = d.ddd x 10^e10
    neg = "-" if v < 0 else ""
    if 1e-3 <= abs(v) < 1e7:
        if e10 >= 0:
            whole = (ds + "0" * (e10 + 1))[: e10 + 1]
            frac = ds[e10 + 1:] or "0"
            return neg + whole + "." + frac
        return neg + "0." + "0" * (-e10 - 1) + ds
    return neg + ds[0] + "." + (ds[1:] or "0") + "E" + str(e10)


def text_lines(text):
    """sc.textFile records."""
    lines = text.split("\n")
    if lines and lines[-1] == "":
        lines.pop()
    return [ln[:-1] if ln.endswith("\r") else ln for ln in lines]


# ---------------------------------------------------------------- shared pieces
def load_results(text):
    """flow_post_lda.scala:105-123 / dns_post_lda.scala:88-106: key -> list of doubles."""
    out = {}
    for line in text_lines(text):
        f = jsplit(line, ",")
        out[f[0]] = [float(v) for v in jsplit(f[1].replace('"', ""), " ")]
    return out


def ecdf(values):
    """compute_ecdf (flow :67-88): [(value, cumulative count / sample size)] ascending."""
    counts = {}
    for v in values:
        counts[v] = counts.get(v, 0) + 1
    n = float(sum(counts.values()))
    run = 0.0
    out = []
    for v in sorted(counts):
        run += counts[v]
        out.append((v, run / n))
    return out


def quantiles(qs, ec):
    """distributed_quantiles (flow :90-103): accumulators start at 0, max over values with ecdf < q."""
    acc = [0.0] * len(qs)
    for v, c in ec:
        for pos, q in enumerate(qs):
            if c < q:
                acc[pos] = max(acc[pos], v)
    return acc


def bin_of(value, cuts):
    b = 0
    for cut in cuts:
        if value > cut:
            b += 1
    return b


QUANT = [0.0, 0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9]
QUINT = [0, 0.2, 0.4, 0.6, 0.8]


# ---------------------------------------------------------------- flow
def flow_post(flow_text, doc_results_text, word_results_text, tol):
    """flow_post_lda.scala:105-248.  Returns dict(cuts, rows (featurised, before scoring),
    indices-free lookups done by string as the script does, scored (output lines))."""
    topics = load_results(doc_results_text)
    words = load_results(word_results_text)
    raw = text_lines(flow_text)
    header = raw[0] if raw else None  # :20-24
    rows = []
    for line in raw:
        if line == header:
            continue
        if len(jsplit(line, ",")) != 27:  # :135
            continue
        row = jsplit(jtrim(line), ",")  # :146
        if len(row) != 27:
            continue  # trimming emptied the last field: the script's row would be short
        num_time = float(row[4]) + float(row[5]) / 60 + float(row[6]) / 3600  # :141
        rows.append(row + [jdouble(num_time)])
    time_cuts = quantiles(QUANT, ecdf([float(r[27]) for r in rows]))  # :150
    ibyt_cuts = quantiles(QUANT, ecdf([float(r[17]) for r in rows]))  # :153
    ipkt_cuts = quantiles(QUINT, ecdf([float(r[16]) for r in rows]))  # :156
    default = [0.05] * 20
    feat, scored = [], []
    for n, row in enumerate(rows):
        # bin_ibyt_ipkt_time (:161-180)
        row = row + [str(bin_of(float(row[17]), ibyt_cuts)), str(bin_of(float(row[16]), ipkt_cuts)),
                     str(bin_of(float(row[27]), time_cuts))]
        # adjust_port (:182-218)
        sip, dip = row[8], row[9]
        dport, sport = float(row[10]), float(row[11])
        ipkt_bin, ibyt_bin, time_bin = float(row[29]), float(row[28]), float(row[30])
        ip_pair = row[9] + " " + row[8]
        if sip < dip:  # `& sip != 0` is String-vs-Int, always true
            ip_pair = row[8] + " " + row[9]
        if (dport <= 1024 or sport <= 1024) and (dport > 1024 or sport > 1024) and min(dport, sport) != 0:
            p_case, word_port = 2, min(dport, sport)
        elif dport > 1024 and sport > 1024:
            p_case, word_port = 3, 333333.0
        elif dport == 0 and sport != 0:
            p_case, word_port = 4, sport
        elif sport == 0 and dport != 0:
            p_case, word_port = 4, dport
        else:
            p_case = 1
            word_port = max(dport, sport) if min(dport, sport) == 0 else 111111.0
        word = jdouble(word_port) + "_" + jdouble(time_bin) + "_" + jdouble(ibyt_bin) + "_" + jdouble(ipkt_bin)
        src_word = dest_word = word
        if p_case == 2 and dport < sport:
            dest_word = "-1_" + dest_word
        elif p_case == 2 and sport < dport:
            src_word = "-1_" + src_word
        elif p_case == 4 and dport == 0:
            src_word = "-1_" + src_word
        elif p_case == 4 and sport == 0:
            dest_word = "-1_" + dest_word
        row = row + [jdouble(word_port), ip_pair, src_word, dest_word]
        feat.append(",".join(row))
        # :227-239
        t1, w1 = topics.get(row[8], default), words.get(row[33], default)
        t2, w2 = topics.get(row[9], default), words.get(row[34], default)
        s1 = s2 = 0.0
        for i in range(20):
            s1 += t1[i] * w1[i]
            s2 += t2[i] * w2[i]
        key = min(s1, s2)
        if key < tol:  # :245
            scored.append((key, n, ",".join(row + [jdouble(s1), jdouble(s2)])))
    scored.sort(key=lambda e: (e[0], e[1]))
    lookups = [(r.split(",")[8] in topics, r.split(",")[9] in topics) for r in feat]
    return dict(cuts=dict(time=time_cuts, ibyt=ibyt_cuts, ipkt=ipkt_cuts), rows=feat,
                scored=[e[2] for e in scored], doc_hits=lookups)


# ---------------------------------------------------------------- dns
COUNTRY_CODES = set(  # dns_post_lda.scala:149, "" and "krd" included
    "ac ad ae af ag ai al am an ao aq ar as at au aw ax az ba bb bd be bf bg bh bi bj bm bn bo bq br bs bt bv bw by "
    "bz ca cc cd cf cg ch ci ck cl cm cn co cr cu cv cw cx cy cz de dj dk dm do dz ec ee eg eh er es et eu fi fj fk "
    "fm fo fr ga gb gd ge gf gg gh gi gl gm gn gp gq gr gs gt gu gw gy hk hm hn hr ht hu id ie il im in io iq ir is "
    "it je jm jo jp ke kg kh ki km kn kp kr krd kw ky kz la lb lc li lk lr ls lt lu lv ly ma mc md me mg mh mk ml mm "
    "mn mo mp mq mr ms mt mu mv mw mx my mz na nc ne nf ng ni nl no np nr nu nz om pa pe pf pg ph pk pl pm pn pr ps "
    "pt pw py qa re ro rs ru rw sa sb sc sd se sg sh si sj sk sl sm sn so sr ss st su sv sx sy sz tc td tf tg th tj "
    "tk tl tm tn to tp tr tt tv tw tz ua ug uk us uy uz va vc ve vg vi vn vu wf ws ye yt za zm zw".split()) | {""}


def extract_subdomain(url):
    """dns_post_lda.scala:154-190."""
    spliturl = jsplit(url, ".")
    numparts = len(spliturl)
    domain = subdomain = "None"
    is_ip = "Unknown"
    if numparts > 2:
        is_ip = "IP" if (spliturl[numparts - 1] == "arpa" and spliturl[numparts - 2] == "in-addr") else "Name"
    if numparts > 2 and is_ip != "IP":
        if spliturl[numparts - 1] in COUNTRY_CODES:
            domain = spliturl[numparts - 3]
            if 1 <= numparts - 3:
                subdomain = ".".join(spliturl[0:numparts - 3])
        else:
            domain = spliturl[numparts - 2]
            if 1 <= numparts - 2:
                subdomain = ".".join(spliturl[0:numparts - 2])
    return [domain, subdomain, str(len(subdomain)) if subdomain != "None" else "0", str(numparts)]


def entropy(v):
    """dns_post_lda.scala:206-212; terms added in ascending character order (unpinned in the script)."""
    total = 0.0
    for ch in sorted(set(v)):
        p = float(v.count(ch)) / len(v)
        total += -p * math.log10(p) / math.log10(2)
    return total


def dns_post(dns_text, top_text, doc_results_text, word_results_text, tol):
    """dns_post_lda.scala:76-330 on comma-joined rows of the eight selected columns (:127)."""
    cols = ["frame_time", "unix_tstamp", "frame_len", "ip_dst", "dns_qry_name", "dns_qry_class", "dns_qry_type",
            "dns_qry_rcode"]
    col = {c: i for i, c in enumerate(cols)}
    top_domains = set(jsplit(jsplit(line, ",")[1], ".")[0] for line in text_lines(top_text))  # :76-81
    topics = load_results(doc_results_text)
    words = load_results(word_results_text)
    data = [f for f in (jsplit(line, ",") for line in text_lines(dns_text)) if len(f) == len(cols)]  # :138
    data = [r + extract_subdomain(r[col["dns_qry_name"]]) for r in data]  # :193
    for name in ("domain", "subdomain", "subdomain.length", "num.periods"):
        col[name] = max(col.values()) + 1
    data = [r + [jdouble(entropy(r[col["subdomain"]]))] for r in data]  # :243
    col["subdomain.entropy"] = max(col.values()) + 1
    time_cuts = quantiles(QUANT, ecdf([float(r[col["unix_tstamp"]]) for r in data]))
    frame_cuts = quantiles(QUANT, ecdf([float(r[col["frame_len"]]) for r in data]))

    def positive(name):
        return [float(r[col[name]]) for r in data if float(r[col[name]]) > 0]

    sub_cuts = quantiles(QUINT, ecdf(positive("subdomain.length")))
    ent_cuts = quantiles(QUINT, ecdf(positive("subdomain.entropy")))
    per_cuts = quantiles(QUINT, ecdf(positive("num.periods")))
    # :272
    data = [r + ["2" if r[col["domain"]] == "intel" else ("1" if r[col["domain"]] in top_domains else "0")]
            for r in data]
    col["top_domain"] = max(col.values()) + 1
    feat, scored = [], []
    default = [0.1] * 20
    for n, r in enumerate(data):
        word = "_".join([r[col["top_domain"]], str(bin_of(float(r[col["frame_len"]]), frame_cuts)),
                         str(bin_of(float(r[col["unix_tstamp"]]), time_cuts)),
                         str(bin_of(float(r[col["subdomain.length"]]), sub_cuts)),
                         str(bin_of(float(r[col["subdomain.entropy"]]), ent_cuts)),
                         str(bin_of(float(r[col["num.periods"]]), per_cuts)),
                         r[col["dns_qry_type"]], r[col["dns_qry_rcode"]]])  # :282-287
        r = r + [word]
        feat.append(",".join(r))
        t = topics.get(r[col["ip_dst"]], default)
        w = words.get(word, default)
        s = 0.0
        for i in range(20):
            s += t[i] * w[i]
        if s < tol:
            scored.append((s, n, ",".join(r + [jdouble(s)])))
    scored.sort(key=lambda e: (e[0], e[1]))
    return dict(cuts=dict(unix_tstamp=time_cuts, frame_len=frame_cuts, subdomain_length=sub_cuts,
                          subdomain_entropy=ent_cuts, num_periods=per_cuts),
                rows=feat, scored=[e[2] for e in scored])


# ---------------------------------------------------------------- the pre-LDA word counts
def flow_word_counts(flow_text):
    """flow_pre_lda.scala:371-382 (same featurisation as the post script, :268-366): "ip,word,count"
    lines, source side then destination side; line order (unspecified in Spark) = first appearance."""
    rows = flow_post(flow_text, "", "", 0.0)["rows"]
    out = []
    for a, b in ((8, 33), (9, 34)):
        counts = {}
        for r in rows:
            f = r.split(",")
            key = f[a] + " " + f[b]
            counts[key] = counts.get(key, 0) + 1
        out += [k.split(" ")[0] + "," + k.split(" ")[1] + "," + str(c) for k, c in counts.items()]
    return out


def dns_word_counts(dns_text, top_text):
    """dns_pre_lda.scala:326 (featurisation :153-323 as in the post script)."""
    rows = dns_post(dns_text, top_text, "", "", 0.0)["rows"]
    counts = {}
    for r in rows:
        f = r.split(",")
        key = f[3] + " " + f[14]
        counts[key] = counts.get(key, 0) + 1
    return [k.split(" ")[0] + "," + k.split(" ")[1] + "," + str(c) for k, c in counts.items()]


# ---------------------------------------------------------------- analyst feedback (pre-LDA scripts)
def flow_feedback_rows(feedback_text, dup):
    """flow_pre_lda.scala:147-263: severity-3 rows of flow_scores.csv as flow rows, dup times each."""
    out = []
    for line in text_lines(feedback_text)[1:]:
        f = jsplit(line, ",")
        if len(f) != 22 or int(f[0]) != 3:
            continue
        hms = jsplit(jsplit(f[1], " ")[1], ":")
        row = ["##"] * 27
        row[4], row[5], row[6] = hms[0], hms[1], hms[2]
        row[8], row[9], row[10], row[11], row[16], row[17] = f[2], f[3], f[4], f[5], f[8], f[9]
        out += [",".join(row)] * dup
    return out


def dns_feedback_rows(feedback_text, dup):
    """dns_pre_lda.scala:70-138: severity-3 rows of dns_scores.csv as the eight selected columns."""
    out = []
    for line in text_lines(feedback_text)[1:]:
        f = jsplit(line, ",")
        if int(jtrim(f[18])) != 3:
            continue
        out += [",".join([f[0], f[23], str(int(jtrim(f[1]))), f[2], f[3], f[4], f[5], f[6]])] * dup
    return out


def with_rows(text, rows):
    if text and not text.endswith("\n"):
        text += "\n"
    return text + "".join(r + "\n" for r in rows)
