ssh test@127.0.0.1 "kill -9 9101"
ssh test@127.0.0.1 "kill -9 9102"
kill -9 8873
