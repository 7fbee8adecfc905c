"""Print the headline fields of a bench.py JSON line.  usage: python scripts/show_bench.py file.json"""
import json, sys
j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value %.0f %s  ms/step %.3f  e2e %.0f  gemm frac %.3f" % (j["value"], j["unit"], j["ms_per_step"], j["e2e"]["value"],
                                                              j["roofline"]["frac"]))
print("kernel ms/step", j.get("kernel_ms_per_step"))
if "latency_bs1" in j:
    print("bs1 p50 device %.3f ms, api %.3f ms" % (j["latency_bs1"]["p50_ms_device"], j["latency_bs1"]["p50_ms_api_host_frame"]))
if "cpu_baseline" in j:
    print("cpu_baseline", j["cpu_baseline"]["value"], j["cpu_baseline"]["unit"])
