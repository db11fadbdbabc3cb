"""Guards for the measurement scripts under scripts/ (CPU only)."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tuning_scripts_do_not_assign_bash_special_variables():
    """`GROUPS=2 python ...` passes nothing: bash keeps GROUPS (and a few others) to itself and silently ignores the
    assignment, so the process runs with its default.  Round 1 and the first sessions of round 2 recorded "two walker
    groups: no gain" that way; the scripts say WGROUPS now.  Keep it that way."""
    special = ("GROUPS", "UID", "EUID", "PPID", "BASHPID", "SECONDS", "RANDOM", "LINENO", "FUNCNAME", "DIRSTACK", "BASH_ARGV0")
    pat = re.compile(r"(?:^|[\s;(])(" + "|".join(special) + r")=")
    bad = []
    for path in sorted(glob.glob(os.path.join(ROOT, "scripts", "*.sh"))):
        for i, line in enumerate(open(path), 1):
            code = line.split("#", 1)[0]
            if pat.search(code):
                bad.append(f"{os.path.relpath(path, ROOT)}:{i}: {line.strip()}")
    assert not bad, "assignment to a bash special variable never reaches the child process:\n" + "\n".join(bad)


def test_profile_case_reads_the_group_count_from_wgroups():
    src = open(os.path.join(ROOT, "scripts", "profile_case.py")).read()
    assert "os.environ['WGROUPS']" in src and "os.environ['GROUPS']" not in src and "os.environ.get('GROUPS'" not in src
