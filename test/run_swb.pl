#!/usr/bin/perl
# Runs the drop-in command line (smithwaterman_b200/swbcli = `sa_opencl`) in list mode for the three alignment types and
# writes one record file per pair and mode: res_swb/<basename of file 1>.res.<mode> -- the layout the reference's
# run_rust_ocl_list.pl produces for check_results.pl.  Usage: perl run_swb.pl [path to swbcli]   (cwd = the data directory)
use strict;
use warnings;
use File::Basename;
use Cwd qw(abs_path);

my $cli = $ARGV[0] // abs_path(dirname(abs_path($0)) . "/../smithwaterman_b200/swbcli");
-x $cli or die "no executable at $cli (build it: python -c 'import __graft_entry__ as g; g.build()')\n";
my $outdir = "res_swb";
-d $outdir or mkdir($outdir) or die "mkdir $outdir: $!";

open(my $lf, "<", "file_list.txt") or die "file_list.txt: $!";
my @names;
while (my $line = <$lf>) {
    my ($first) = split(/[\t ]/, $line);
    next unless defined $first && length $first;
    push @names, basename($first);
}
close($lf);

foreach my $mode ("local", "glocal", "global") {
    my $text = `"$cli" -$mode -list file_list.txt`;
    $? == 0 or die "$cli -$mode -list failed ($?)\n";
    my @records = grep { length } split(/(?=^#score)/m, $text);
    scalar(@records) == scalar(@names) or die "$mode: " . scalar(@records) . " records for " . scalar(@names) . " pairs\n";
    for my $k (0 .. $#names) {
        open(my $o, ">", "$outdir/$names[$k].res.$mode") or die $!;
        print $o $records[$k];
        close($o);
    }
}
print "wrote " . (3 * scalar(@names)) . " records to $outdir/\n";
