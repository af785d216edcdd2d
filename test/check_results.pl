#!/usr/bin/perl
# The reference's acceptance rules (test/check_results.pl of yamule/smithwaterman) applied to res_swb/:
#   1. an output record with its gaps removed spells the two input sequences;
#   2. the aligned strings equal the EMBOSS result of the same pair: water (res<N>.dat) for local after the unaligned
#      overhangs are trimmed, needle without end weights (needle_glocal_res<N>.dat) for glocal, needle -endweight
#      (needle_res<N>.dat) for global.
# A difference is fatal unless a sequence holds one of J U Z B O X (letters EMBOSS scores differently).
# Prints "OK" and "Checked <n> results".  cwd = the data directory.
use strict;
use warnings;

sub slurp_record {          # (>s1 block, >s2 block) of a record or FASTA-like file, whitespace removed
    my ($path) = @_;
    open(my $in, "<", $path) or die "$path: $!";
    my ($which, @seq) = (0, "", "");
    while (my $l = <$in>) {
        if ($l =~ />s1/) { $which = 1; next; }
        if ($l =~ />s2/) { $which = 2; next; }
        $seq[$which - 1] .= $l if $which;
    }
    close($in);
    s/\s//g for @seq;
    return @seq;
}

sub emboss_strings {        # the s1 / s2 rows of an srspair report, concatenated
    my ($path) = @_;
    open(my $in, "<", $path) or die "$path: $!";
    my ($a, $b) = ("", "");
    while (my $l = <$in>) {
        $a .= $1 if $l =~ /^\s*s1\s+[0-9]+\s*(\S+)/;
        $b .= $1 if $l =~ /^\s*s2\s+[0-9]+\s*(\S+)/;
    }
    close($in);
    return ($a, $b);
}

sub trim_terminal {         # drop leading / trailing columns until both rows hold a letter
    my @x = split(//, $_[0]);
    my @y = split(//, $_[1]);
    my ($lo, $hi) = (0, $#x);
    $lo++ while $lo <= $hi && ($x[$lo] eq "-" || $y[$lo] eq "-");
    $hi-- while $hi >= $lo && ($x[$hi] eq "-" || $y[$hi] eq "-");
    return (join("", @x[$lo .. $hi]), join("", @y[$lo .. $hi]));
}

my %golden_prefix = (local => "res", glocal => "needle_glocal_res", global => "needle_res");
opendir(my $dh, "res_swb") or die "res_swb: $! (run run_swb.pl first)";
my @files = sort grep { /\.res\.(local|glocal|global)$/ } readdir($dh);
closedir($dh);
my $checked = 0;
foreach my $f (@files) {
    my ($base, $mode) = $f =~ /^(.*)\.res\.(\w+)$/;
    my ($id) = $base =~ /seq([0-9]+)/ or die "unexpected name $f";
    my ($r1, $r2) = slurp_record("res_swb/$f");
    (my $file2 = $base) =~ s/\.1\.fas/.2.fas/;
    my ($in1) = slurp_record("inputs/$base");
    my (undef, $in2) = slurp_record("inputs/$file2");
    my $tolerated = ($r1 . $r2) =~ /[JUZBOX]/;
    (my $p1 = $r1) =~ s/[^A-Z]//g;
    (my $p2 = $r2) =~ s/[^A-Z]//g;
    if (($p1 ne $in1 || $p2 ne $in2) && !$tolerated) {
        die "$f: the record does not spell its input sequences\n$p1\n$in1\n$p2\n$in2\n";
    }
    my @got = $mode eq "local" ? trim_terminal($r1, $r2) : ($r1, $r2);
    my @want = emboss_strings("emboss_results/$golden_prefix{$mode}$id.dat");
    if (($got[0] ne $want[0] || $got[1] ne $want[1]) && !$tolerated) {
        die "$f: differs from EMBOSS\n$got[0]\n$want[0]\n$got[1]\n$want[1]\n";
    }
    $checked++;
}
print "OK\n";
print "Checked $checked results\n";
