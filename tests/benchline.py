import json,sys
d=json.load(open(sys.argv[1]));print(sys.argv[1].split('/')[-1], "val %.0f e2e %.0f convTF %.0f"%(d["value"], d["e2e"]["value"], d["roofline"]["achieved"]), d["stages_ms"])
