"""write_example_conf: the `-i` template (src/Conf_file.py:15-104).

Same sections, option names and example values as the reference template (so the documented smoke test on
test/dataset runs unchanged); the explanatory comments are this project's own wording.
"""

TEMPLATE = """\
# ------------------------------------------------------------------------------------------------
#  QUADE configuration file (INI syntax: `option : value`, '#' starts a comment line)
#  Keep every section and option; edit the values only.  Paths must not contain blanks.
# ------------------------------------------------------------------------------------------------

[quality]
# Qualities must be Phred+33 (Illumina 1.8+).
# A pair passes only if EVERY base of its sample index has at least this phred score; 0 = no filter. (integer 0-40)
minimal_qual : 25

[fastq]
# One list of files per read, chunk by chunk and in the same order in every list.
# Single index: seq_R1 = R1, seq_R2 = R3, index_R1 = R2.
# Dual index:   seq_R1 = R1, seq_R2 = R4, index_R1 = R2, index_R2 = R3.
seq_R1 :   ../dataset/C1_R1.fastq.gz  ../dataset/C2_R1.fastq.gz  ../dataset/C3_R1.fastq.gz
seq_R2 :   ../dataset/C1_R4.fastq.gz  ../dataset/C2_R4.fastq.gz  ../dataset/C3_R4.fastq.gz
index_R1 : ../dataset/C1_R2.fastq.gz  ../dataset/C2_R2.fastq.gz  ../dataset/C3_R2.fastq.gz
index_R2 : ../dataset/C1_R3.fastq.gz  ../dataset/C2_R3.fastq.gz  ../dataset/C3_R3.fastq.gz

[index]
# index1 is always used.  Switch the second sample index and the molecular indexes on or off. (boolean)
index2 : True
molecular1 : True
molecular2 : True

# 1-based, inclusive positions of each sub-index inside its index read; only read for the sub-indexes
# switched on above.  Sub-indexes may overlap. (integer)
index1_start : 1
index1_end : 4
index2_start : 1
index2_end : 4
molecular1_start : 4
molecular1_end : 6
molecular2_start : 4
molecular2_end : 6

[output]
# Which classes of pairs are written to fastq.gz files; the report counts all of them anyway. (boolean)
write_pass : True
write_fail : True
write_undetermined : True

# One [sampleN] section per sample, any number of them, reported in this order.
# name: unique, used as file prefix.  index1_seq / index2_seq: upper-case A, C, G, T or N; the fused index
# (index1_seq followed by index2_seq) must be unique.  index2_seq is read only when index2 is True.
[sample1]
name : S1
index1_seq : ACAG
index2_seq : ACAG

[sample2]
name : S2
index1_seq : CTTG
index2_seq : CTTG
"""


def write_example_conf():
    with open("Quade_conf_file.txt", "w") as fp:
        fp.write(TEMPLATE)
