"""Run Hadoop Pipes children without a JVM: write a command file, start the child with
`hadoop.pipes.command.file` set, parse `<file>.out` (HadoopPipes.cc:1130-1136)."""
import os
import subprocess
import tempfile
from concurrent.futures import ThreadPoolExecutor

from hgpumap import wire


def run_child(binary, downlink: wire.Downlink, extra_env=None, timeout=120, args=()):
    tmp = tempfile.mkdtemp(prefix="pipes_")
    cmd = os.path.join(tmp, "cmd")
    downlink.write(cmd)
    env = {"hadoop.pipes.command.file": cmd, "PATH": os.environ.get("PATH", "")}
    for key in ("LD_LIBRARY_PATH", "CUDA_VISIBLE_DEVICES"):
        if key in os.environ:
            env[key] = os.environ[key]
    if extra_env:
        env.update(extra_env)
    proc = subprocess.run([binary, *args], env=env, capture_output=True, timeout=timeout)
    out = cmd + ".out"
    data = open(out, "rb").read() if os.path.exists(out) else b""
    return proc, data, tmp


def run_children(jobs, max_workers=8):
    """jobs: list of (binary, downlink, extra_env).  The reference child's ping thread
    sleeps 5 s before it lets runTask return (HadoopPipes.cc:1049), so run them together."""
    with ThreadPoolExecutor(max_workers=max_workers) as pool:
        futs = [pool.submit(run_child, *job) for job in jobs]
        return [f.result() for f in futs]


TESTPIPES_INPUT = [  # src/test/org/apache/hadoop/mapred/pipes/TestPipes.java:138-147
    "Alice was beginning to get very tired of sitting by her",
    "sister on the bank, and of having nothing to do: once",
    "or twice she had peeped into the book her sister was",
    "reading, but it had no pictures or conversations in",
    "it, `and what is the use of a book,' thought Alice",
    "`without pictures or conversation?'",
]

# TestPipes.java:126-133 (fixedPartitionOutput[0]): wordcount-part, everything to reduce 0
FIXED_PARTITION_OUTPUT = (
    "Alice\t2\n`and\t1\n`without\t1\na\t1\nand\t1\nbank,\t1\nbeginning\t1\n"
    "book\t1\nbook,'\t1\nbut\t1\nby\t1\nconversation?'\t1\nconversations\t1\n"
    "do:\t1\nget\t1\nhad\t2\nhaving\t1\nher\t2\nin\t1\ninto\t1\nis\t1\n"
    "it\t1\nit,\t1\nno\t1\nnothing\t1\nof\t3\non\t1\nonce\t1\nor\t3\n"
    "peeped\t1\npictures\t2\nreading,\t1\nshe\t1\nsister\t2\nsitting\t1\n"
    "the\t3\nthought\t1\ntired\t1\nto\t2\ntwice\t1\nuse\t1\n"
    "very\t1\nwas\t2\nwhat\t1\n"
)

# TestPipes.java:114-124 (noSortOutput): wordcount-simple with 0 reduces, one file per map;
# file order there is by split (lines 5-6, 1-2, 3-4), here re-ordered to input order.
NO_SORT_OUTPUT_IN_INPUT_ORDER = (
    "Alice\t1\nwas\t1\nbeginning\t1\nto\t1\nget\t1\nvery\t1\ntired\t1\n"
    "of\t1\nsitting\t1\nby\t1\nher\t1\nsister\t1\non\t1\nthe\t1\nbank,\t1\n"
    "and\t1\nof\t1\nhaving\t1\nnothing\t1\nto\t1\ndo:\t1\nonce\t1\n"
    "or\t1\ntwice\t1\nshe\t1\nhad\t1\npeeped\t1\ninto\t1\nthe\t1\nbook\t1\n"
    "her\t1\nsister\t1\nwas\t1\nreading,\t1\nbut\t1\nit\t1\nhad\t1\nno\t1\n"
    "pictures\t1\nor\t1\nconversations\t1\nin\t1\n"
    "it,\t1\n`and\t1\nwhat\t1\nis\t1\nthe\t1\nuse\t1\nof\t1\na\t1\n"
    "book,'\t1\nthought\t1\nAlice\t1\n`without\t1\npictures\t1\nor\t1\n"
    "conversation?'\t1\n"
)


def wordcount_downlink(num_reduces: int) -> wire.Downlink:
    d = (wire.Downlink().authenticate().start()
         .set_job_conf({"io.sort.mb": "1", "mapred.task.partition": "0"})
         .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.Text")
         .run_map(b"ignored-split", num_reduces, True))
    off = 0
    for line in TESTPIPES_INPUT:
        d.map_item(off.to_bytes(8, "big"), line)  # LongWritable key = 8 bytes big-endian
        off += len(line) + 1
    return d.close()


def kmeans_downlink(points, centroid_path, k, d, fmt="binary", num_reduces=1, per_record=1, io_sort_mb=100,
                    extra_conf=None):
    """Downlink for a K-means map task with piped input.  `per_record` points per MAP_ITEM."""
    import numpy as np

    conf = {"kmeans.k": str(k), "kmeans.d": str(d), "kmeans.centroids": centroid_path,
            "kmeans.input.format": fmt, "io.sort.mb": str(io_sort_mb),
            "hadoop.pipes.java.recordreader": "true"}  # piped input (pipes/Submitter.java:128-139)
    conf.update(extra_conf or {})
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf)
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.BytesWritable")
          .run_map(b"ignored-split", num_reduces, True))
    pts = np.ascontiguousarray(points, dtype="<f4")
    n = pts.shape[0]
    if fmt == "binary":
        rec = per_record * d * 4
        whole = (n // per_record) * per_record
        if whole and rec < 128:
            vals = pts[:whole].view(np.uint8).reshape(-1, rec)
            keys = (np.arange(vals.shape[0], dtype=">u8") * rec).view(np.uint8).reshape(-1, 8)
            dl.map_items_raw(wire.frame_fixed_map_items(keys, vals))
        else:
            for i in range(0, whole, per_record):
                dl.map_item((i * d * 4).to_bytes(8, "big"), pts[i:i + per_record].tobytes())
        if whole < n:
            dl.map_item((whole * d * 4).to_bytes(8, "big"), pts[whole:].tobytes())
    elif fmt == "text":
        for i in range(n):
            dl.map_item(i.to_bytes(8, "big"), " ".join(repr(float(v)) for v in pts[i]))
    else:
        raise ValueError(fmt)
    return dl.close()


# ---- socket mode: play the Java task (pipes/Application.java:130-215) --------------------------

def java_digest(password: bytes, msg: bytes) -> bytes:
    """base64(HMAC-SHA1(password, msg)) as SecureShuffleUtils.hashFromString / the child's
    createDigest (HadoopPipes.cc:421-448) compute it."""
    import base64
    import hashlib
    import hmac

    return base64.b64encode(hmac.new(password, msg, hashlib.sha1).digest())


def run_child_over_socket(binary, downlink_after_auth: wire.Downlink, password=b"job-token-secret",
                          challenge=b"1234567890123456", bad_digest=False, timeout=60, args=()):
    """Listen on a loopback port, start the child with hadoop.pipes.command.port and the
    password file, authenticate, send the downlink, collect the uplink until EOF."""
    import socket

    tmp = tempfile.mkdtemp(prefix="pipes_sock_")
    secret = os.path.join(tmp, "jobTokenPassword")
    with open(secret, "wb") as f:
        f.write(password)
    os.chmod(secret, 0o400)
    server = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    server.bind(("127.0.0.1", 0))
    server.listen(4)
    port = server.getsockname()[1]
    env = {"hadoop.pipes.command.port": str(port), "hadoop.pipes.shared.secret.location": secret,
           "TMPDIR": tmp, "PATH": os.environ.get("PATH", "")}
    for key in ("LD_LIBRARY_PATH", "CUDA_VISIBLE_DEVICES"):
        if key in os.environ:
            env[key] = os.environ[key]
    proc = subprocess.Popen([binary, *args], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    server.settimeout(timeout)
    conn, _ = server.accept()
    conn.settimeout(timeout)
    digest = java_digest(password, challenge)
    if bad_digest:
        digest = b"A" + digest[1:]
    expected = java_digest(password, digest)
    auth = wire.Downlink().authenticate(digest.decode(), challenge.decode()).to_bytes()
    conn.sendall(auth + downlink_after_auth.to_bytes())
    chunks = []
    try:
        while True:
            data = conn.recv(1 << 16)
            if not data:
                break
            chunks.append(data)
    except (socket.timeout, ConnectionResetError):
        pass
    out, err = proc.communicate(timeout=timeout)
    conn.close()
    server.close()
    return proc.returncode, b"".join(chunks), err, expected
