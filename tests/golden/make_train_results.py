"""Golden vectors for the monitoring parser: runs the reference's own TensorResultView.read_train_results
(/root/reference/apps/runtime/views.py:44-72; the method body is executed from the reference's source text because the
module imports Django, which is not installed) on result.txt samples and stores (text, iter, returned JSON).
Run HERE:  python tests/golden/make_train_results.py  ->  tests/golden/train_results.json"""
import json
import os
import re
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
src = open("/root/reference/apps/runtime/views.py").read()
body = re.search(r"    def read_train_results\(self,filepath,iter\):.*?return json_data\n", src, re.S).group(0)
ns = {}
exec("import json\nclass V:\n" + body, ns)

# the sample response of API.md:454-511 has this shape: step lines every 100 steps, then final_accuracy
full = "".join("step:%d,accuracy:%f,duration:%f\n" % (s, a, d) for s, a, d in
               [(0, 0.08, 0.0), (100, 0.16, 0.141), (200, 0.3, 0.138), (300, 0.56, 0.151)]) + "final_accuracy:%f\n\n" % 0.44
cases = []
for text in (full, full[:full.index("final")], full.split("\n")[0] + "\n", "", full.replace("final_accuracy:0.440000\n\n", "final_accuracy:0.44")):
    for it in (100, 300, 1000, "200"):
        with tempfile.NamedTemporaryFile("w", delete=False) as f:
            f.write(text)
        cases.append({"text": text, "iter": it, "json": ns["V"]().read_train_results(f.name, it)})
        os.unlink(f.name)
json.dump(cases, open(os.path.join(HERE, "train_results.json"), "w"), indent=1)
print(len(cases), "cases")
