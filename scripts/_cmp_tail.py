import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1], {k:d["hyperpartition"][k] for k in ("cluster_ms_each","cluster_ms","groups","modularity","wall_s")})
