"""Build helper: generate OINK's style_*.h in a scratch copy of oink/ (python tools/oink_styles.py <scratch dir>).
OINK does this with a Python-2 script (oink/Make.py); it is run unmodified except for the mechanical py2 -> py3 fixes
SURVEY.md 8c lists (the `commands` module, `print >>fp, x`), applied to a copy inside the scratch directory."""
import os
import re
import subprocess
import sys


def generate(scratch: str) -> None:
    txt = open(os.path.join(scratch, "Make.py")).read()
    txt = txt.replace("import sys,os,glob,commands,re", "import sys,os,glob,re\nimport subprocess as commands")
    txt = re.sub(r"print >>(\w+),(.*)", r"print(\2, file=\1)", txt)
    with open(os.path.join(scratch, "Make3.py"), "w") as f:
        f.write(txt)
    subprocess.check_call([sys.executable, "Make3.py"], cwd=scratch)


if __name__ == "__main__":
    generate(sys.argv[1])
