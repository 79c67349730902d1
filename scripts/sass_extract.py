"""Cut one function's section out of `nvdisasm -g -c` output.  Usage: sass_extract.py ALL.sass MANGLED_NAME > fn.sass"""
import sys
lines = open(sys.argv[1]).read().split('\n')
name = sys.argv[2]
start = None
for i, l in enumerate(lines):
    if l.strip() == '.section\t.text.' + name + ',"ax",@progbits':
        start = i
        break
assert start is not None, "function not found"
end = len(lines)
for j in range(start + 3, len(lines)):
    if lines[j].lstrip().startswith('.section'):
        end = j
        break
print('\n'.join(lines[start:end]))
