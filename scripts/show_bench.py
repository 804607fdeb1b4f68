import json,sys
for f in sys.argv[1:]:
    j=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, round(j["value"]), round(j["ms_per_step"],3), round(j["roofline"]["kernel_ms"],3), j["config"]["exchange"][:12], j["config"]["exchange_check"])
