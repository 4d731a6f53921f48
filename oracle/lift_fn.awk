# lift_fn.awk — streams ONE function definition out of a reference source file (test infrastructure, oracle/Makefile).
# usage: awk -v pat='<regex matching the line that carries the function name>' -f lift_fn.awk file.cc
# Prints the line before the match when the match is a continuation line (starts with white space: the return type sits
# on the line above), then every line up to and including the first line that starts with '}' (or that ends with an
# empty body '{}').
{
  if (!p && $0 ~ pat) {
    p = 1
    if ($0 ~ /^[ \t]/ && prev != "") {
      print prev
    }
  }
  if (p) {
    print
    if ($0 ~ /^}/ || $0 ~ /\{\}[ \t]*$/) {
      exit
    }
  }
  prev = $0
}
